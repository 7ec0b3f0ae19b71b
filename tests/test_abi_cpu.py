"""The C-ABI shared library loads on a machine without a GPU and exports every symbol include/rtb200.h declares;
compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(rtb):
    lib = C.CDLL(rtb._abi.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"librtb200.so does not export {n}"
    assert set(names) == set(rtb._abi.SYMBOLS), "the ctypes binding and the header disagree"


def test_struct_layouts_match_header(rtb, tmp_path):
    """sizeof of every struct as the C compiler sees include/rtb200.h == sizeof of the ctypes mirror."""
    import subprocess
    A = rtb._abi
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "rtb200.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(rtb200_material),sizeof(rtb200_sphere),sizeof(rtb200_scene),sizeof(rtb200_camera),'
                   'sizeof(rtb200_params),sizeof(rtb200_stats));return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(t) for t in (A.Material, A.Sphere, A.Scene, A.Camera, A.Params, A.Stats)]


def test_camera_init_is_host_only(rtb):
    cam = rtb.Camera.default(400, 225)     # pure host arithmetic: works without a device
    assert cam.image_width() == 400
    with pytest.raises(rtb.Rtb200Error):
        rtb.Camera(0, 10, 10.0, 30.0, (0, 0, 1), (0, 0, 0), (0, 1, 0), 0.0)


def test_no_cpu_fallback(rtb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the no-device error path is not reachable")
    with pytest.raises(rtb.Rtb200Error) as e:
        rtb.CudaRenderer(device=0)
    assert e.value.code == rtb._abi.ERR_CUDA and "no CPU fallback" in e.value.message


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "rust-ray-tracing_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower() or f == "rt_device.cuh", f"{f} mentions the oracle"


def test_null_context_is_an_error_on_every_device_entry_point(rtb):
    """Argument validation runs before any CUDA call: a null context is RTB200_ERR_INVALID on a box without a GPU too
    (nothing aborts across the ABI).  Covers rtb200_trace_device_static, the entry point without a reference counterpart."""
    import ctypes as C
    lib, A = rtb._abi.lib, rtb._abi
    cam = rtb.Camera.default(16, 9)
    p = A.Params(5, 2, 1, 0, 2, 0, 9, 0, 1)
    used = C.c_uint64(123)
    assert lib.rtb200_trace_device_static(None, C.byref(cam.abi), C.byref(p), None, None, C.byref(used)) == A.ERR_INVALID
    assert lib.rtb200_trace_device(None, C.byref(cam.abi), C.byref(p), None, None) == A.ERR_INVALID
    assert lib.rtb200_render(None, C.byref(cam.abi), C.byref(p), None, None, None) == A.ERR_INVALID
    assert lib.rtb200_device_free(None, None) == A.ERR_INVALID and lib.rtb200_ipc_close(None, None) == A.ERR_INVALID
    assert b"null" in lib.rtb200_last_error() or b"ctx" in lib.rtb200_last_error()
    assert A.STATIC_SCRATCH_BYTES == 32 << 20
    header = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    assert "#define RTB200_STATIC_SCRATCH_BYTES %d" % A.STATIC_SCRATCH_BYTES in header
