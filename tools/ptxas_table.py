#!/usr/bin/env python
"""Registers / spills / stack per kernel from the ptxas -v log of the build (build/obj/ptxas_api.log)."""
import re
import subprocess
import sys


def parse(path):
    rows, name = [], None
    for line in open(path):
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("rt::", "")
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and name:
            stack, st, ld = map(int, m.groups())
            continue_row = [name, stack, st, ld]
            rows.append(continue_row)
            continue
        m = re.search(r"Used (\d+) registers", line)
        if m and rows and len(rows[-1]) == 4:
            rows[-1].append(int(m.group(1)))
    return rows


if __name__ == "__main__":
    path = sys.argv[1] if len(sys.argv) > 1 else "build/obj/ptxas_api.log"
    print(f"{'kernel':60s} {'regs':>5s} {'stack':>6s} {'spill st':>8s} {'spill ld':>8s}")
    for r in parse(path):
        print(f"{r[0]:60s} {r[4] if len(r) > 4 else -1:5d} {r[1]:6d} {r[2]:8d} {r[3]:8d}")
