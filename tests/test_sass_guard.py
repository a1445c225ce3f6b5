"""Static checks on the SASS of the shipped library (CPU only; needs cuobjdump from the CUDA toolkit).

* ptxas 12.9 was seen to put a predicated CREDUX (warp reduce to a uniform register) of gather_kernel under an UNRELATED live
  predicate, while the instructions consuming its result used the right one (round 2: positions-model gains of buses with a
  non-trivial passthrough came out as stale uniform-register contents).  The source now keeps every warp reduction
  unconditional; this test fails the build if a predicated one comes back.
* The window staging engine the build claims (TMA bulk copy, SASS UBLKCP + SYNCS, or cp.async, SASS LDGSTS) is the one in
  the binary, and the packed FP32x2 arithmetic is there.
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "libzl_b200", "libzl_b200.so")


def sass():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe) or not os.path.exists(SO):
        pytest.skip("cuobjdump or the built library is missing")
    return subprocess.run([exe, "-sass", SO], capture_output=True, text=True, check=True).stdout


def test_no_predicated_warp_reduction_in_the_render_kernels():
    fn, bad = None, []
    for line in sass().splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
        if fn and ("gather_kernel" in fn or "plan_kernel" in fn or "walk_kernel" in fn) and re.search(r"@!?U?P\d+\s+(C?REDUX)", line):
            bad.append((fn, line.strip()[:90]))
    assert not bad, bad[:4]


def test_staging_engine_and_packed_arithmetic_are_in_the_binary():
    text = sass()
    src = open(os.path.join(ROOT, "libzl_b200", "csrc", "zl_render.cu")).read()
    tma_default = re.search(r"#define ZL_TMA (\d)", src).group(1) == "1"
    if tma_default:
        assert "UBLKCP" in text and "SYNCS.PHASECHK" in text, "TMA bulk-copy staging is the default but UBLKCP is not in the SASS"
    else:
        assert "LDGSTS" in text
    assert "FFMA2" in text and "FMUL2" in text and "FADD2" in text
