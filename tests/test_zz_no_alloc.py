"""omg_debug_ddiv_static (include/omg_gpu.h): the allocation-free variant of omg_debug_ddiv.

No counterpart in the reference; the entry point was added for callers that cannot
call cudaMalloc.  The device staging of the operands and the quotients sits in a file-scope __device__ array of
3 x OMG_DDIV_STATIC_MAX doubles (csrc/omg_kernels.cu), whose address the library takes with cudaGetSymbolAddress.

The file name puts this test last in pytest's collection order.  It goes through the C-ABI only, twice in a row (the second
call finds the first call's values in the shared array), and compares with omg_debug_ddiv: the same kernel on the same
values, so the comparison is bitwise -- the tolerance of the other division test (tests/test_gpu_steps.py)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX = 262144      # OMG_DDIV_STATIC_MAX


def operands(seed, n):
    rng = np.random.default_rng(seed)
    a = rng.uniform(-4.0, 4.0, n) * 10.0 ** rng.integers(-3, 4, n)
    b = rng.uniform(0.25, 4.0, n) * 10.0 ** rng.integers(-3, 4, n)
    wide = slice(0, n // 3)                       # full exponent range: candidates for the slow path of the division
    a[wide] = rng.uniform(-2, 2, n // 3) * 2.0 ** rng.integers(-1070, 1023, n // 3)
    b[wide] = rng.uniform(1, 2, n // 3) * 2.0 ** rng.integers(-1070, 1023, n // 3) * rng.choice([-1.0, 1.0], n // 3)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, 1.0, 3.0])
    k = len(special)
    a[n - k * k:] = np.repeat(special, k)
    b[n - k * k:] = np.tile(special, k)
    return a, b


def ptr(x):
    return x.ctypes.data_as(C.c_void_p)


def same_bits(x, y):
    return (x.view(np.uint64) == y.view(np.uint64)) | (np.isnan(x) & np.isnan(y))


def test_static_scratch_variant_matches_the_allocating_entry_point(product):
    G, _ = product.libs()
    err = C.create_string_buffer(256)
    # two calls in a row with different sizes and values: the second one must not see anything of the first
    for seed, n, batched in ((11, MAX, 1), (12, 6 * 20000 + 5, 1), (13, 6 * 1000 + 1, 0)):
        a, b = operands(seed, n)
        want = np.full(n, -1.0)
        got = np.full(n, -2.0)
        assert G.omg_debug_ddiv(0, ptr(a), ptr(b), ptr(want), n, batched) == 0
        rc = G.omg_debug_ddiv_static(0, ptr(a), ptr(b), ptr(got), n, batched, err, len(err))
        assert rc == 0, err.value
        assert err.value == b""
        ok = same_bits(got, want)
        assert ok.all(), (seed, n, a[~ok][:5], b[~ok][:5], got[~ok][:5], want[~ok][:5])
        with np.errstate(all="ignore"):
            assert same_bits(got, a / b).all()


def test_static_scratch_variant_rejects_sizes_beyond_its_array(product):
    G, _ = product.libs()
    err = C.create_string_buffer(256)
    n = MAX + 1
    a, b = np.ones(n), np.ones(n)
    out = np.zeros(n)
    assert G.omg_debug_ddiv_static(0, ptr(a), ptr(b), ptr(out), n, 1, err, len(err)) == 2
    assert b"exceeds the static scratch" in err.value and str(MAX).encode() in err.value
    assert not out.any()                                    # nothing was written
    assert G.omg_debug_ddiv_static(0, ptr(a), ptr(b), ptr(out), n, 1, None, 0) == 2      # no message buffer
    assert G.omg_debug_ddiv_static(0, ptr(a), ptr(b), ptr(out), 0, 1, err, len(err)) == 0   # empty input: nothing to do
    assert G.omg_debug_ddiv_static(0, None, ptr(b), ptr(out), 8, 1, err, len(err)) == 3
    # and the entry point still works after the refusals
    a2, b2 = operands(5, 6000)
    got, want = np.empty(6000), np.empty(6000)
    assert G.omg_debug_ddiv_static(0, ptr(a2), ptr(b2), ptr(got), 6000, 1, err, len(err)) == 0, err.value
    assert G.omg_debug_ddiv(0, ptr(a2), ptr(b2), ptr(want), 6000, 1) == 0
    assert same_bits(got, want).all()
