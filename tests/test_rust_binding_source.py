"""rust/ffi.rs cannot be compiled here (no Rust toolchain): check at least that it stays in step with include/rtgpu.h —
every #[repr(C)] struct has the header's fields in the header's order, every extern fn exists in the header, the
constants agree, and the flatten() visitor covers every object / material / texture kind of the ABI."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _c_structs():
    text = open(os.path.join(ROOT, "include", "rtgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", text, flags=re.S):
        names = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for piece in decl.split(","):
                ident = re.findall(r"([A-Za-z_]\w*)\s*(?:\[[^\]]*\])*\s*$", piece.strip())
                names.append(ident[0])
        out[m.group(1)] = names
    return out, text


def _rust_structs():
    text = open(os.path.join(ROOT, "rust", "ffi.rs")).read()
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^)]*\)\]\s*)?pub struct (\w+) \{(.*?)\n\}", text, flags=re.S):
        out[m.group(1)] = re.findall(r"pub (\w+):", m.group(2))
    return out, text


def test_rust_repr_c_structs_match_the_header_field_for_field():
    c, _ = _c_structs()
    r, _ = _rust_structs()
    for name in ("RtgpuObject", "RtgpuMesh", "RtgpuMaterial", "RtgpuTexture", "RtgpuImage", "RtgpuNoise", "RtgpuSceneDesc",
                 "RtgpuCameraDesc", "RtgpuRenderParams"):
        assert name in c and name in r, name
        assert c[name] == r[name], (name, c[name], r[name])


def test_rust_extern_block_and_constants_match_the_header():
    _, ctext = _c_structs()
    _, rtext = _rust_structs()
    fns = re.findall(r"pub fn (rtgpu_\w+)\(", rtext)
    assert len(fns) >= 10
    for f in fns:
        assert re.search(r"\b%s\s*\(" % f, ctext), f"{f} is not declared in include/rtgpu.h"
    enums = dict(re.findall(r"\b(RTGPU_(?:OBJ|MAT|TEX|IMG)_\w+) = (\d+)", ctext))
    for cname, val in enums.items():
        rname = cname.replace("RTGPU_", "")
        m = re.search(r"pub const %s: i32 = (\d+);" % rname, rtext)
        assert m and m.group(1) == val, (cname, val)


def test_flatten_visitor_covers_every_kind_of_the_abi():
    src = open(os.path.join(ROOT, "rust", "flatten.rs")).read() + open(os.path.join(ROOT, "rust", "flatten_materials.rs")).read()
    _, rtext = _rust_structs()
    for prefix in ("OBJ_", "MAT_", "TEX_"):
        kinds = re.findall(r"pub const (%s\w+): i32" % prefix, rtext)
        assert len(kinds) == {"OBJ_": 10, "MAT_": 7, "TEX_": 8}[prefix]
        for k in kinds:
            assert re.search(r"\b%s\b" % k, src), f"no flatten() emits {k}"
    assert src.count("pub fn flatten_impl") == 10 + 7 + 8 + 1   # objects, materials, textures, PerlinNoise3D
