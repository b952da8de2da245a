"""Per-kernel GPU parity: every building block of the cycle (smooth_boxes, update_coarse = residual +
restriction + coarse rhs, correct_children = correction + prolongation, max_residual_lvl) run on its own on
seeded random states and compared bitwise with the oracle, level by level."""
import numpy as np
import pytest

import problems as P
from compare import assert_state_equal

pytestmark = pytest.mark.gpu


def randomize(mg, seed, variables=(1, 2, 3, 4)):
    rng = np.random.default_rng(seed)
    for lvl, id_ in P.all_ids(mg):
        cc = mg.cc(id_)
        for iv in variables:
            cc[iv - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal(cc[0, 1:-1, 1:-1, 1:-1].shape)


def prepare(oracle, product, n, box, op=P.mg_laplacian, bc="callback", lam=0.0, seed=7):
    orc, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, n, box, operator_type=op, bc=bc, lam=lam)
    for m in (orc, mg):
        randomize(m, seed)
    # make the oracle's stored ghost cells consistent with the interiors: old(:) = phi(:) of a first random
    # field (ghosts included), then a second random field in phi
    orc.mg_fill_ghost_cells(P.mg_iphi)
    for lvl, id_ in P.all_ids(orc):
        c_o, c_g = orc.cc(id_), mg.cc(id_)
        c_o[P.mg_iold - 1] = c_o[P.mg_iphi - 1]
        c_g[P.mg_iold - 1] = c_g[P.mg_iphi - 1]
    randomize(orc, seed + 1, variables=(1,))
    randomize(mg, seed + 1, variables=(1,))
    orc.mg_fill_ghost_cells(P.mg_iphi)
    return orc, mg


CASES = [(32, 8, P.mg_laplacian, "callback", 0.0), ((48, 48, 48), 8, P.mg_laplacian, "callback", 0.0),
         ((32, 64, 96), 16, P.mg_helmholtz, "mixed", 1.0), (32, 8, P.mg_vlaplacian, "mixed", 0.0)]


@pytest.mark.parametrize("n,box,op,bc,lam", CASES)
def test_smooth_boxes_every_level(oracle, product, n, box, op, bc, lam):
    orc, mg = prepare(oracle, product, n, box, op, bc, lam)
    for lvl in range(orc.highest_lvl, orc.lowest_lvl - 1, -1):
        orc.smooth_boxes(lvl, 2)
        mg.smooth_boxes(lvl, 2)
        assert_state_equal(orc, mg, variables=(1,), levels=[lvl])


@pytest.mark.parametrize("n,box,op,bc,lam", CASES)
def test_update_coarse_every_level(oracle, product, n, box, op, bc, lam):
    orc, mg = prepare(oracle, product, n, box, op, bc, lam)
    for lvl in range(orc.highest_lvl, orc.lowest_lvl, -1):
        orc.update_coarse(lvl)
        mg.update_coarse(lvl)
        assert_state_equal(orc, mg, variables=(4,), ghosts_of=(), levels=[lvl])           # fine residual
        assert_state_equal(orc, mg, variables=(1, 2, 3, 4), ghosts_of=(1, 3), levels=[lvl - 1])


@pytest.mark.parametrize("n,box,op,bc,lam", CASES)
def test_correct_children_every_level(oracle, product, n, box, op, bc, lam):
    orc, mg = prepare(oracle, product, n, box, op, bc, lam)
    for lvl in range(orc.lowest_lvl, orc.highest_lvl):
        orc.correct_children(lvl)
        orc.mg_fill_ghost_cells_lvl(lvl + 1, P.mg_iphi)
        mg.correct_children(lvl)
        assert_state_equal(orc, mg, variables=(1,), levels=[lvl + 1])
        assert_state_equal(orc, mg, variables=(4,), ghosts_of=(), levels=[lvl])          # res = phi - old on parents


@pytest.mark.parametrize("n,box,op,bc,lam", CASES)
def test_max_residual_every_level(oracle, product, n, box, op, bc, lam):
    orc, mg = prepare(oracle, product, n, box, op, bc, lam)
    for lvl in range(orc.lowest_lvl, orc.highest_lvl + 1):
        assert orc.max_residual_lvl(lvl) == mg.max_residual_lvl(lvl)
        assert_state_equal(orc, mg, variables=(4,), ghosts_of=(), levels=[lvl])


def test_timers(product):
    """mg_timers_show: the reference's seven timers (src/m_multigrid.f90:72-81); totals always, phases in profile mode."""
    mg, _ = P.uniform_problem(product.Mg, 32, 8)
    mg.mg_fas_fmg(False)
    mg.mg_fas_vcycle()
    t = mg.mg_timers()
    assert list(t) == ["mg total V-cycle", "mg total FMG cycle", "mg smoother", "mg smoother g.c.", "mg coarse",
                       "mg correct", "mg update coarse"]
    assert t["mg total V-cycle"] > 0 and t["mg total FMG cycle"] > 0 and t["mg smoother"] == 0
    mg.set_option("profile", 1)
    mg.mg_fas_vcycle()
    mg.set_option("profile", 0)
    t2 = mg.mg_timers()
    assert t2["mg smoother"] > 0 and t2["mg coarse"] > 0 and t2["mg correct"] > 0 and t2["mg update coarse"] > 0
    assert t2["mg total V-cycle"] > t["mg total V-cycle"]


def test_batched_division_is_ieee_division(product):
    """ddiv_batch (omg_device.cuh), the branch-free division of the variable-coefficient kernels, against the
    compiler's `/` on the device and numpy's on the host: bitwise, including the operands nvcc's fast path rejects"""
    import ctypes as C
    G, _ = product.libs()
    rng = np.random.default_rng(7)
    n = 6 * 200000
    a = rng.uniform(-4.0, 4.0, n) * 10.0 ** rng.integers(-3, 4, n)
    b = rng.uniform(0.25, 4.0, n) * 10.0 ** rng.integers(-3, 4, n)
    wide = slice(0, n // 3)                       # full exponent range: slow-path candidates mixed into batches
    a[wide] = rng.uniform(-2, 2, n // 3) * 2.0 ** rng.integers(-1070, 1023, n // 3)
    b[wide] = rng.uniform(1, 2, n // 3) * 2.0 ** rng.integers(-1070, 1023, n // 3) * rng.choice([-1.0, 1.0], n // 3)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, 1.0, 3.0])
    k = len(special)
    a[n - k * k:] = np.repeat(special, k)
    b[n - k * k:] = np.tile(special, k)
    out = [np.empty(n), np.empty(n)]
    for batched in (0, 1):
        rc = G.omg_debug_ddiv(0, a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p),
                              out[batched].ctypes.data_as(C.c_void_p), n, batched)
        assert rc == 0
    with np.errstate(all="ignore"):
        ref = a / b
    for got in out:
        same = (got.view(np.uint64) == ref.view(np.uint64)) | (np.isnan(got) & np.isnan(ref))
        assert same.all(), (a[~same][:5], b[~same][:5], got[~same][:5], ref[~same][:5])
