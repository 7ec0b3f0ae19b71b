"""integration/rust: the Rust side of the drop-in boundary as files (build.rs, src/rtb200_sys.rs, src/cuda_renderer.rs) and
integration/reference.patch.  No Rust toolchain exists here, so what CAN be checked is checked:
  * rtb200_sys.rs is exactly what tools/gen_rust_sys.py generates from include/rtb200.h (no drift);
  * every #[repr(C)] struct in it, laid out by the C rules from its Rust field types, has the size and field offsets gcc
    reports for the header;
  * every function the header declares is declared in the extern block with the same number of parameters, and
    cuda_renderer.rs only calls functions that exist;
  * reference.patch applies cleanly to the reference tree (when /root/reference is present) and only touches the files
    INTEGRATION.md names."""
import os
import re
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SYS_RS = os.path.join(ROOT, "integration", "rust", "src", "rtb200_sys.rs")
RENDERER_RS = os.path.join(ROOT, "integration", "rust", "src", "cuda_renderer.rs")
HEADER = os.path.join(ROOT, "include", "rtb200.h")

RUST_SCALAR = {"i32": (4, 4), "u32": (4, 4), "i64": (8, 8), "u64": (8, 8), "f64": (8, 8), "f32": (4, 4), "u8": (1, 1), "usize": (8, 8), "c_int": (4, 4)}


def test_sys_rs_is_generated_from_the_header():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_rust_sys.py"), "--check"])
    assert r.returncode == 0, "integration/rust/src/rtb200_sys.rs is stale: run python tools/gen_rust_sys.py"


def _rust_structs():
    text = open(SYS_RS).read()
    out = {}
    for name, body in re.findall(r"#\[repr\(C\)\]\s*#\[derive\([^)]*\)\]\s*pub struct (\w+) \{(.*?)\n\}", text, flags=re.S):
        fields = []
        for fname, ftype in re.findall(r"pub (\w+): ([^,\n]+),", body):
            fields.append((fname, ftype.strip()))
        out[name] = fields
    return out


def _layout(fields):
    off, align_max, offs = 0, 1, {}
    for name, t in fields:
        m = re.match(r"\[(\w+); (\d+)\]", t)
        if m:
            size, align = RUST_SCALAR[m.group(1)]
            size *= int(m.group(2))
        elif t.startswith("*"):
            size, align = 8, 8
        else:
            size, align = RUST_SCALAR[t]
        off = (off + align - 1) // align * align
        offs[name] = off
        off += size
        align_max = max(align_max, align)
    return (off + align_max - 1) // align_max * align_max, offs


def test_rust_struct_layouts_match_gcc(tmp_path):
    structs = _rust_structs()
    assert set(structs) == {"rtb200_material", "rtb200_sphere", "rtb200_scene", "rtb200_camera", "rtb200_params", "rtb200_stats"}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "rtb200.h"', "int main(void){"]
    for s, fields in structs.items():
        lines.append(f'printf("{s} %zu", sizeof({s}));')
        for f, _ in fields:
            lines.append(f'printf(" {f}=%zu", offsetof({s}, {f}));')
        lines.append('printf("\\n");')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    got = {}
    for ln in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines():
        parts = ln.split()
        got[parts[0]] = (int(parts[1]), {kv.split("=")[0]: int(kv.split("=")[1]) for kv in parts[2:]})
    for s, fields in structs.items():
        size, offs = _layout(fields)
        assert (size, offs) == got[s], s


def test_extern_block_covers_the_header():
    hdr = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    c_funcs = {}
    for name, args in re.findall(r"\b(rtb200_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = " ".join(args.split())
        c_funcs[name] = 0 if args in ("", "void") else args.count(",") + 1
    rs = open(SYS_RS).read()
    rs_funcs = {n: (0 if not a.strip() else a.count(":")) for n, a in re.findall(r"pub fn (rtb200_\w+)\((.*?)\)(?: ->|;)", rs)}
    assert rs_funcs == c_funcs
    used = set(re.findall(r"\b(rtb200_[a-z0-9_]+)\s*\(", open(RENDERER_RS).read()))
    assert used and used <= set(c_funcs), used - set(c_funcs)
    # the constants cuda_renderer.rs names exist
    for c in set(re.findall(r"\b(RTB200_[A-Z0-9_]+)\b", open(RENDERER_RS).read())) - {"RTB200_GPUS"}:
        assert re.search(rf"pub const {c}: i32", rs), c


def test_reference_patch_applies(tmp_path):
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "src")):
        pytest.skip("the reference tree is not present on this machine")
    patch = os.path.join(ROOT, "integration", "reference.patch")
    touched = set(re.findall(r"^\+\+\+ b/(\S+)", open(patch).read(), flags=re.M))
    assert touched == {"src/color.rs", "src/main.rs", "src/materials.rs", "src/objects.rs", "src/ray_tracing.rs"}
    work = tmp_path / "ref"
    shutil.copytree(ref, work, ignore=shutil.ignore_patterns(".git"))
    r = subprocess.run(["patch", "-p1", "--dry-run", "-i", patch], cwd=work, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
