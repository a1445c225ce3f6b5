"""The C-ABI library loads and exports every symbol include/zl_b200.h declares (no compute calls: CPU only)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "zl_b200.h")


def header_functions():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"ZL_FN\((\w+)\)\s*\(", src)))


def test_header_declares_the_hot_path():
    names = header_functions()
    for must in ("ctx_create", "clip_create_from_pcm", "command", "render", "read_bus", "read_master", "levels_tick",
                 "thumb_build", "thumb_pixels", "thumb_raster", "passthrough_process", "levels_scan"):
        assert must in names
    assert len(names) >= 38


def test_product_library_exports_every_declared_symbol():
    import libzl_b200
    from libzl_b200 import _capi
    assert os.path.exists(libzl_b200.LIB_PATH), "build with: python -c 'import __graft_entry__ as g; g.build()'"
    dll = C.CDLL(libzl_b200.LIB_PATH)
    missing = [n for n in header_functions() if not hasattr(dll, "zl_" + n)]
    assert not missing, missing
    assert sorted(_capi.SIGNATURES) == header_functions(), "ctypes bindings out of sync with the header"
    lib = libzl_b200.load()
    assert lib.prefix == "zl_"


def test_oracle_exports_the_same_abi(oracle):
    missing = [n for n in header_functions() if not hasattr(oracle.dll, "zlo_" + n)]
    assert not missing, missing


def test_product_has_no_cpu_fallback():
    """Without a CUDA device ctx_create must fail loudly with ZL_ERR_NODEVICE."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import libzl_b200
    from libzl_b200 import _capi as K
    from libzl_b200.engine import Engine, ZlError
    lib = libzl_b200.load()
    cfg = K.zl_config()
    assert lib.config_default(C.byref(cfg)) == K.ZL_OK
    ctx = C.c_void_p()
    assert lib.ctx_create(C.byref(cfg), C.byref(ctx)) == K.ZL_ERR_NODEVICE
    with pytest.raises(ZlError):
        Engine()


def test_product_never_references_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "libzl_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "from oracle" not in src and "import oracle" not in src, f
    out = os.popen(f"ldd {os.path.join(ROOT, 'libzl_b200', 'libzl_b200.so')}").read()
    assert "oracle" not in out
