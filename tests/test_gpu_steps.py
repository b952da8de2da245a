"""Per-kernel GPU parity: every building block of the cycle (smooth_boxes, update_coarse = residual +
restriction + coarse rhs, correct_children = correction + prolongation, max_residual_lvl) run on its own on
seeded random states and compared bitwise with the oracle, level by level."""
import os

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


@pytest.mark.parametrize("amr", [False, True])
def test_copy_to_and_from_tree(product, amr):
    """omg_copy_to_tree / omg_copy_from_tree (mg_copy_to_tree, mg_copy_from_tree[_gc] of
    coupling_amrvac/mod_multigrid_coupling.t:133-267): host buffer <-> device tree, leaves or whole levels, with or
    without the ghost layer, against the host cc arrays moved by mg_gpu_push / mg_gpu_pull"""
    if amr:
        mg, _ = P.refinement_problem(product.Mg, 3, 8, 16, gpu_resident=True)
    else:
        mg, _ = P.uniform_problem(product.Mg, 32, 8, bc="dirichlet0", gpu_resident=True)
    rng = np.random.default_rng(3)
    lvls = range(mg.lowest_lvl, mg.highest_lvl + 1)
    for leaves_only in (True, False):
        boxes = [(lvl, id_) for lvl in lvls for id_ in (mg.leaves(lvl) if leaves_only else mg.ids(lvl))]
        # interior cells in: every other cell of the device copy must keep its previous value
        for _, id_ in P.all_ids(mg):
            mg.cc(id_)[P.mg_irhs - 1] = -7.0
        mg.mg_gpu_push(P.mg_irhs)
        parts = [rng.standard_normal(mg.box_size_lvl(lvl) ** 3) for lvl, _ in boxes]
        assert mg.copy_size(leaves_only, False) == sum(p.size for p in parts)
        mg.mg_copy_to_tree(P.mg_irhs, np.concatenate(parts), leaves_only=leaves_only)
        mg.mg_gpu_pull(P.mg_irhs, False)
        touched = {id_ for _, id_ in boxes}
        for (lvl, id_), part in zip(boxes, parts):
            nc = mg.box_size_lvl(lvl)
            assert np.array_equal(mg.cc(id_)[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1], part.reshape(nc, nc, nc))
        for _, id_ in P.all_ids(mg):
            if id_ not in touched:
                assert np.all(mg.cc(id_)[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] == -7.0)
        # out again, interior only and with the ghost layer phi's boundary conditions / neighbours define
        got = mg.mg_copy_from_tree(P.mg_irhs, leaves_only=leaves_only)
        assert np.array_equal(got, np.concatenate(parts))
        for _, id_ in P.all_ids(mg):
            cc = mg.cc(id_)
            cc[P.mg_iphi - 1] = rng.standard_normal(cc[P.mg_iphi - 1].shape)
        mg.mg_gpu_push(P.mg_iphi)
        got = mg.mg_copy_from_tree(P.mg_iphi, leaves_only=leaves_only, with_ghosts=True)
        mg.mg_gpu_pull(P.mg_iphi, True)
        want = np.concatenate([mg.cc(id_)[P.mg_iphi - 1].ravel() for _, id_ in boxes])
        from compare import face_slices
        off = 0
        for lvl, id_ in boxes:          # compare interiors and the six faces (edges / corners of the shell are not defined)
            n2 = mg.box_size_lvl(lvl) + 2
            a = got[off:off + n2 ** 3].reshape(n2, n2, n2)
            b = want[off:off + n2 ** 3].reshape(n2, n2, n2)
            assert np.array_equal(a[1:-1, 1:-1, 1:-1], b[1:-1, 1:-1, 1:-1])
            for sl in face_slices(n2 - 2):
                assert np.array_equal(a[sl], b[sl])
            off += n2 ** 3


@pytest.mark.parametrize("op,bc,lam", [(P.mg_laplacian, "callback", 0.0), (P.mg_helmholtz, "callback", 2.5),
                                       (P.mg_vlaplacian, "mixed", 0.0), (P.mg_vhelmholtz, "neumann0", 1.0),
                                       (P.mg_ahelmholtz, "mixed", 0.5)])
def test_tail_fast_path_all_variants(oracle, product, op, bc, lam):
    orc, _ = P.uniform_problem(oracle.OracleMg, (64, 32, 32), 16, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, (64, 32, 32), 16, operator_type=op, bc=bc, lam=lam)
    mg.set_option("tail_simple", 2)
    assert P.run_cycles(orc, 3, True) == P.run_cycles(mg, 3, True)
    assert P.run_cycles(orc, 2, False) == P.run_cycles(mg, 2, False)
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("op,lam", [(P.mg_laplacian, 0.0), (P.mg_vhelmholtz, 1.5)])
@pytest.mark.parametrize("graph", [0, 1])
def test_pdl_launches_change_nothing(oracle, product, op, lam, graph):
    orc, _ = P.uniform_problem(oracle.OracleMg, (64, 64, 32), 16, operator_type=op, bc="mixed", lam=lam)
    mg, _ = P.uniform_problem(product.Mg, (64, 64, 32), 16, operator_type=op, bc="mixed", lam=lam)
    mg.set_option("tail_cells", 512)      # several levels on the TMA-staged kernels, back to back
    mg.set_option("use_graph", graph)
    mg.set_option("pdl", 1)
    assert P.run_cycles(orc, 3, True) == P.run_cycles(mg, 3, True)
    assert P.run_cycles(orc, 3, False) == P.run_cycles(mg, 3, False)
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("warp", [1, 0])
@pytest.mark.parametrize("op,bc,lam", [(P.mg_laplacian, "callback", 0.0), (P.mg_laplacian, "neumann0", 0.0), (P.mg_helmholtz, "neumann0", 1.0),
                                       (P.mg_vlaplacian, "mixed", 0.0), (P.mg_vhelmholtz, "neumann0", 1.0),
                                       (P.mg_ahelmholtz, "mixed", 0.5)])
def test_warp_coarse_solve(oracle, product, op, bc, lam, warp):
    """cubic grids end in ONE 2^3 box: its coarse solve (src/m_multigrid.f90:202-208) runs in one warp with the eight cells in
    registers (omg_tail.cuh warp_coarse_solve; option warp_coarse = 0: as interpreter operations) -- residual histories, the
    full state and the iteration count of the coarse loop as the oracle's, for problems whose coarse loop is short (Dirichlet)
    and long (the constant mode of a Neumann problem: tens of iterations, or the limit max_coarse_cycles)"""
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 16, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, 32, 16, operator_type=op, bc=bc, lam=lam)
    mg.set_option("warp_coarse", warp)
    for m in (orc, mg):
        m.set_int("max_coarse_cycles", 30)      # the Neumann problems reach the limit (the oracle's coarse loop is slow per iteration)
    assert P.run_cycles(orc, 2, True) == P.run_cycles(mg, 2, True)
    assert P.run_cycles(orc, 1, False) == P.run_cycles(mg, 1, False)
    assert_state_equal(orc, mg)
    assert 1 <= mg.stat("coarse_iters") <= 30
