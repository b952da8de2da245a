"""Multi-rank GPU parity on ONE device: n emulated ranks (one handle + one host thread each, omg_comm_init_local)
run the reference's Morton partition (mg_load_balance) with the static exchange plans of omg_plan.cpp -- ghost
faces, restriction blocks, coarse correction patches, scalar all-reduce -- and must reproduce the oracle
bitwise (results are independent of the number of ranks).  The NCCL transport carries the very
same buffers; it is exercised by bench.py --gpus N."""
import functools

import pytest

import problems as P
from compare import assert_state_equal

pytestmark = pytest.mark.gpu


def group_cls(product, n_cpu):
    return functools.partial(product.MgGroup, n_cpu)


def check(h_o, h_g):
    assert h_o == h_g, "residual histories differ: %r vs %r" % (h_o, h_g)


@pytest.mark.parametrize("n_cpu", [2, 3, 8])
@pytest.mark.parametrize("fmg", [False, True])
def test_uniform_ranks(oracle, product, n_cpu, fmg):
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8, n_cpu=n_cpu)
    mg, _ = P.uniform_problem(group_cls(product, n_cpu), 32, 8)
    check(P.run_cycles(orc, 3, fmg), P.run_cycles(mg, 3, fmg))
    assert_state_equal(orc, mg)
    assert mg.stat("exchanges") > 0


@pytest.mark.parametrize("lb", ["default", "simple", "parents"])
def test_fast_kernels_with_remote_faces(oracle, product, lb):
    """box 16: the TMA-staged kernels run on levels whose partition faces arrive through the plane pool"""
    orc, _ = P.uniform_problem(oracle.OracleMg, (64, 32, 32), 16, n_cpu=4, load_balance=lb)
    mg, _ = P.uniform_problem(group_cls(product, 4), (64, 32, 32), 16, load_balance=lb)
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    # the halo exchange of the smoother ran inside the smoother kernels (peer-mapped puts + flags), the other ghost
    # fills as put kernels
    assert mg.stat("fused_smooths") > 0 and mg.stat("puts") > 0
    # ... and the prolongation kernels that precede such half-sweeps put their corrected faces themselves
    assert mg.stat("fused_prolongs") > 0
    # ... and every level kernel of the box-16 levels was a TMA-staged one (omg_fast.cu), on every rank
    assert mg.stat("fast_launches") > 0 and mg.stat("general_cells") <= 0.2 * mg.stat("fast_cells")


@pytest.mark.parametrize("opts", [{"fused": 0}, {"fused": 0, "overlap": 0}, {"p2p": 0}, {"p2p": 0, "overlap": 0}])
def test_exchange_transports(oracle, product, opts):
    """every transport of the ghost exchange gives the same bits: in-kernel puts (default), stand-alone put kernels with
    and without the boundary-first overlap, and packed send/recv messages (the ncclSend/ncclRecv path)"""
    orc, _ = P.uniform_problem(oracle.OracleMg, (64, 32, 32), 16, n_cpu=4)
    mg, _ = P.uniform_problem(group_cls(product, 4), (64, 32, 32), 16)
    for k, v in opts.items():
        mg.set_option(k, v)
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    assert mg.stat("fused_smooths") == 0
    assert (mg.stat("puts") > 0) == (opts.get("p2p", 1) == 1)


@pytest.mark.parametrize("n_cpu,dims", [(2, (64, 32, 32)), (8, (64, 64, 64))])
def test_fused_prolongation_put_on_and_off(oracle, product, n_cpu, dims):
    """the ghost exchange after correct_children (src/m_multigrid.f90:219-222) issued from inside the prolongation kernel
    (option fused_prolong, default) against the stand-alone put kernels: same bits, and the stand-alone exchange phases it
    replaces are gone"""
    orc, _ = P.uniform_problem(oracle.OracleMg, dims, 16, n_cpu=n_cpu)
    h_o = P.run_cycles(orc, 2, True) + P.run_cycles(orc, 2, False)
    puts = {}
    for on in (1, 0):
        mg, _ = P.uniform_problem(group_cls(product, n_cpu), dims, 16)
        mg.set_option("fused_prolong", on)
        check(h_o, P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False))
        assert_state_equal(orc, mg)
        assert (mg.stat("fused_prolongs") > 0) == (on == 1)
        puts[on] = mg.stat("puts")
    assert puts[1] < puts[0]


@pytest.mark.parametrize("op,bc,lam", [(P.mg_helmholtz, "neumann0", 1.0), (P.mg_vlaplacian, "mixed", 0.0)])
def test_operators_ranks(oracle, product, op, bc, lam):
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8, operator_type=op, bc=bc, lam=lam, n_cpu=2)
    mg, _ = P.uniform_problem(group_cls(product, 2), 32, 8, operator_type=op, bc=bc, lam=lam)
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("n_cpu", [2, 5])
def test_refinement_ranks(oracle, product, n_cpu):
    """AMR: refinement-boundary faces whose coarse neighbour lives on another rank are interpolated by the
    coarse owner and sent (buffer_for_fine_nb, src/m_ghost_cells.f90:385-422)."""
    def build(cls, **kw):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, **kw)
        for nb in range(1, 7):
            mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.25)
        mg.mg_set_methods()
        P.build_amr_tree(mg, 3, [32] * 3, 8, [1.0 / 32] * 3)
        for lvl, id_ in P.all_ids(mg):
            mg.cc(id_)[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = 1.0 + 0.1 * lvl
        return mg
    orc, mg = build(oracle.OracleMg, n_cpu=n_cpu), build(group_cls(product, n_cpu))
    check(P.run_cycles(orc, 3, True), P.run_cycles(mg, 3, True))
    check(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)


def test_periodic_ranks(oracle, product):
    """periodic neighbours across ranks + subtract_mean (all-reduce SUM); tolerance as in the single-rank test"""
    import numpy as np

    def build(cls, **kw):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, **kw)
        mg.mg_set_methods()
        mg.mg_build_rectangle([32] * 3, 8, [1.0 / 32] * 3, [0.0] * 3, [True] * 3)
        mg.mg_load_balance()
        mg.mg_allocate_storage()
        for lvl, id_ in P.all_ids(mg):
            X, Y, Z = P.cell_coords(mg, id_, lvl)
            mg.cc(id_)[P.mg_irhs - 1] = np.sin(2 * np.pi * X) * np.cos(4 * np.pi * Y) + 0.3 * np.sin(2 * np.pi * Z) + 0.05
        return mg
    orc, mg = build(oracle.OracleMg, n_cpu=4), build(group_cls(product, 4))
    rhs_norm = 1.4
    for fmg in (True, False):
        for a, b in zip(P.run_cycles(orc, 3, fmg), P.run_cycles(mg, 3, fmg)):
            assert abs(a - b) <= max(1e-10 * abs(a), 1e-12 * rhs_norm)


def test_steps_ranks(oracle, product):
    """building blocks on their own across 4 ranks"""
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8, n_cpu=4)
    mg, _ = P.uniform_problem(group_cls(product, 4), 32, 8)
    orc.mg_fill_ghost_cells(P.mg_iphi)     # the device derives phi's ghost cells; make the oracle's stored ones consistent
    for lvl in range(orc.highest_lvl, orc.lowest_lvl, -1):
        orc.smooth_boxes(lvl, 1)
        mg.smooth_boxes(lvl, 1)
        orc.update_coarse(lvl)
        mg.update_coarse(lvl)
    assert_state_equal(orc, mg, ghosts_of=(1,))
    for lvl in range(orc.lowest_lvl, orc.highest_lvl):
        orc.correct_children(lvl)
        orc.mg_fill_ghost_cells_lvl(lvl + 1, P.mg_iphi)
        mg.correct_children(lvl)
    assert_state_equal(orc, mg, variables=(1, 4), ghosts_of=(1,))
    for lvl in range(orc.lowest_lvl, orc.highest_lvl + 1):
        assert orc.max_residual_lvl(lvl) == mg.max_residual_lvl(lvl)


def test_coarse_level_collapse(oracle, product):
    """coarse levels gathered onto one rank (host-mirror option coarse_collapse_boxes): same numbers"""
    def make(*a, **k):
        g = product.MgGroup(4, *a, **k)
        g.set_int("coarse_collapse_boxes", 16)
        return g
    orc = P.performance_problem(oracle.OracleMg, 64, 8, n_cpu=4)
    mg = P.performance_problem(make, 64, 8)
    owners = {lvl: sorted({mg.box_info(i)["rank"] for i in mg.ids(lvl)}) for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1)}
    assert len(owners[mg.highest_lvl]) == 4 and len(owners[mg.highest_lvl - 2]) == 1, owners
    check(P.run_cycles(orc, 3, False), P.run_cycles(mg, 3, False))
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("n_cpu", [2, 4])
def test_gs_smoother_ranks(oracle, product, n_cpu):
    """mg_smoother_gs across ranks: with Gauss-Seidel the result depends on the box size but not on the number of
    ranks; the ghost snapshot taken before each sweep reads remote faces from the exchanged plane pool."""
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8, smoother=P.mg_smoother_gs, n_cpu=n_cpu)
    mg, _ = P.uniform_problem(group_cls(product, n_cpu), 32, 8, smoother=P.mg_smoother_gs)
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("op,lam", [(P.mg_vlaplacian, 0.0), (P.mg_ahelmholtz, 0.75)])
def test_fast_variable_coefficient_ranks(oracle, product, op, lam):
    """vlaplacian / ahelmholtz at box 16 on 4 ranks: fast kernels (with the in-kernel halo exchange) with phi faces from the
    plane pool and eps faces from the stored pool"""
    orc, _ = P.uniform_problem(oracle.OracleMg, (64, 32, 32), 16, operator_type=op, bc="mixed", lam=lam, n_cpu=4)
    mg, _ = P.uniform_problem(group_cls(product, 4), (64, 32, 32), 16, operator_type=op, bc="mixed", lam=lam)
    check(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    assert mg.stat("fast_launches") > 0 and mg.stat("general_cells") <= 0.2 * mg.stat("fast_cells")
