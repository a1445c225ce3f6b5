"""zl_levels_scan_static: the allocation-free variant of the standalone peak / sum-of-squares scan (no counterpart in the
reference).  Its scratch is a file-scope static __device__ array whose device
address the library looks up with cudaGetSymbolAddress; here it is called through the C ABI and compared with the existing
entry point zl_levels_scan (and the oracle) on calls in a row, at sizes up to the stated maximum, and refused beyond it.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from libzl_b200 import _capi as K
from libzl_b200.engine import Engine

MAX_FRAMES = 262144          # ZL_STATIC_SCAN_MAX_FRAMES, include/zl_b200.h


def _scan(e, fn, l, r):
    pa, pb, sa, sb = C.c_int32(-1), C.c_int32(-1), C.c_double(-1.0), C.c_double(-1.0)
    lp = l.ctypes.data_as(C.POINTER(C.c_float))
    rp = r.ctypes.data_as(C.POINTER(C.c_float))
    rc = fn(e.ctx, lp, rp, l.size, C.byref(pa), C.byref(pb), C.byref(sa), C.byref(sb))
    return rc, (pa.value, pb.value, sa.value, sb.value)


def _same(a, b):
    assert a[:2] == b[:2]                                              # integer peaks: bit-exact
    for x, y in zip(a[2:], b[2:]):                                     # f64 sums, order-free: the tolerance of test_parity_gpu.py
        assert abs(x - y) <= 1e-9 * max(abs(y), 1e-300)


@pytest.mark.gpu
def test_static_scratch_scan_equals_the_allocating_entry_point(product, oracle):
    rng = np.random.default_rng(2024)
    eg, eo = Engine(lib=product), Engine(lib=oracle)
    before = eg.launch_count()
    # calls in a row on the shared array: a long one first, then shorter ones that must not see what it left behind
    for n in (100003, 4096, 4096, 255, 1, 0, MAX_FRAMES, 777):
        l = rng.uniform(-1.5, 1.5, n).astype(np.float32)
        r = rng.uniform(-0.25, 0.25, n).astype(np.float32)
        rc, got = _scan(eg, product.levels_scan_static, l, r)
        assert rc == K.ZL_OK, eg.last_error() if hasattr(eg, "last_error") else rc
        rc2, ref = _scan(eg, product.levels_scan, l, r)
        assert rc2 == K.ZL_OK
        _same(got, ref)
        _same(got, _scan(eo, oracle.levels_scan, l, r)[1])
        if n:
            assert got[0] > got[1] > 0 and got[2] > got[3] > 0.0       # a real scan, not zeros
        else:
            assert got == (0, 0, 0.0, 0.0)
    assert eg.launch_count() >= before + 14                            # the kernel ran for both entry points, every non-empty call
    # the same input twice in a row gives the same integers (the accumulators are cleared, not carried over)
    l = rng.uniform(-1, 1, 5000).astype(np.float32)
    a = _scan(eg, product.levels_scan_static, l, -l)[1]
    b = _scan(eg, product.levels_scan_static, l, -l)[1]
    assert a[:2] == b[:2] and a[0] == a[1]
    _same(a, b)
    # beyond the fixed scratch: refused with a clear error, on the product and on the oracle's mirror of the boundary
    big = np.zeros(MAX_FRAMES + 1, np.float32)
    rc, _ = _scan(eg, product.levels_scan_static, big, big)
    assert rc == K.ZL_ERR_INVALID
    assert b"ZL_STATIC_SCAN_MAX_FRAMES" in product.last_error(eg.ctx)
    assert _scan(eo, oracle.levels_scan_static, big, big)[0] == K.ZL_ERR_INVALID
    # ... and the context still works afterwards
    _same(_scan(eg, product.levels_scan_static, l, -l)[1], a)
    eg.close(); eo.close()


@pytest.mark.gpu
def test_static_scratch_scan_from_a_second_context(product):
    """Two contexts on the device share the one array (serialised here): the per-device address is looked up once."""
    rng = np.random.default_rng(7)
    l = rng.uniform(-1, 1, 30000).astype(np.float32)
    r = rng.uniform(-1, 1, 30000).astype(np.float32)
    e1, e2 = Engine(lib=product), Engine(lib=product)
    a = _scan(e1, product.levels_scan_static, l, r)[1]
    b = _scan(e2, product.levels_scan_static, l, r)[1]
    c = _scan(e2, product.levels_scan, l, r)[1]
    _same(a, b); _same(a, c)
    e1.close(); e2.close()
