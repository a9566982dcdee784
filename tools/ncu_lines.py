#!/usr/bin/env python
"""tools/ncu_lines.py <report.ncu-rep> <mangled-kernel-prefix> [launch-index] — per-source-line instruction and stall-sample shares.

ncu's CSV source page carries per-SASS-instruction counters but no line numbers; nvdisasm --print-line-info of the
cubin in rust_raytracer_b200/librtgpu.so carries the line of every instruction. Both list the kernel's instructions
in address order, so they are zipped. The library must be the build the report was taken from.
"""
import collections, csv, io, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, prefix = sys.argv[1], sys.argv[2]
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "rust_raytracer_b200", "librtgpu.so")], cwd=tmp, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if "sm_100" in f][0]
    dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.split("\n")
    start = [i for i, l in enumerate(dis) if l.startswith(".text." + prefix)][0]
    end = [i for i, l in enumerate(dis) if i > start and l.lstrip().startswith(".section")][0]
    cur, lines = None, []
    for l in dis[start:end]:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        elif re.match(r"\s*/\*[0-9a-f]{4,6}\*/", l):
            lines.append(cur)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = [i for i, x in enumerate(rows) if x and x[0] == "Address"]
    h = rows[hdr[which]]
    body = [x for x in rows[hdr[which] + 1: hdr[which + 1] - 1 if len(hdr) > which + 1 else None] if len(x) >= len(h)]
    assert len(body) == len(lines), (len(body), len(lines), "library and report are different builds")
    iE, iN = h.index("Instructions Executed"), h.index("# Samples")
    ex, sm = collections.Counter(), collections.Counter()
    for loc, x in zip(lines, body):
        ex[loc] += int(x[iE] or 0); sm[loc] += int(x[iN] or 0)
    te, ts = sum(ex.values()), sum(sm.values())
    byf_e, byf_s = collections.Counter(), collections.Counter()
    for k, v in ex.items():
        byf_e[k[0] if k else None] += v
    for k, v in sm.items():
        byf_s[k[0] if k else None] += v
    print(f"{len(body)} instructions, {te} executed, {ts} samples")
    for f, v in byf_e.most_common(12):
        print(f"  {str(f):28s} instr {v / te * 100:5.1f}%  samples {byf_s[f] / ts * 100:5.1f}%")
    print("top lines by executed instructions / by samples:")
    for (k, v), (k2, v2) in zip(ex.most_common(25), sm.most_common(25)):
        print(f"  {str(k):34s} {v / te * 100:5.2f}% (samples {sm[k] / ts * 100:5.2f}%)    |  {str(k2):34s} samples {v2 / ts * 100:5.2f}% (instr {ex[k2] / te * 100:5.2f}%)")


if __name__ == "__main__":
    main()
