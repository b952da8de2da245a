"""GPU parity: the CUDA path (through the C ABI and the host mirror) against the CPU oracle on the same
inputs.  The product is compiled with -fmad=false and uses the reference's expression order, so for GSRB
every number must agree BITWISE with the oracle (-ffp-contract=off); the north-star tolerances (1e-10
relative on per-cycle residuals, 1e-12 * |phi| on the solution) are asserted as well for the record."""
import numpy as np
import pytest

import problems as P
from compare import assert_state_equal, state_diff

pytestmark = pytest.mark.gpu

RES_RTOL = 1e-10     # north-star: per-cycle residual norms within 1e-10 relative
SOL_RTOL = 1e-12     # north-star: final solution max-abs difference within 1e-12 * solution norm


def check_histories(h_o, h_g):
    assert len(h_o) == len(h_g)
    for a, b in zip(h_o, h_g):
        assert abs(a - b) <= RES_RTOL * abs(a), (h_o, h_g)
    assert h_o == h_g, "residual histories differ in the last bits: %r vs %r" % (h_o, h_g)


def assert_fast_path(mg, general_ok=False):
    """the TMA-staged kernels of omg_fast.cu must have run (omg_stat counters): a silent fall-back to the general
    one-CTA-per-box path would keep the numbers right and the performance claim wrong"""
    f, g = mg.stat("fast_launches"), mg.stat("general_launches")
    assert f > 0, "no TMA-staged kernel ran (fast %d, general %d)" % (f, g)
    if general_ok == "small":
        # transitions to a level of another box size (below first_normal_lvl) run on the general path: a few small
        # levels; weighted by cells the TMA-staged kernels must still do nearly all of the work
        fc, gc = mg.stat("fast_cells"), mg.stat("general_cells")
        assert gc <= 0.2 * fc, "general-path share too large (fast %d cells, general %d cells)" % (fc, gc)
    elif not general_ok:
        assert g == 0, "level kernels fell back to the general path (fast %d, general %d)" % (f, g)


def check_tolerance(orc, mg):
    d = state_diff(orc, mg, variables=(1,))["phi"]
    assert d[0] <= SOL_RTOL * d[1]


@pytest.mark.parametrize("fmg", [False, True])
@pytest.mark.parametrize("n,box", [(32, 8), (64, 16)])
def test_uniform_grid_laplacian(oracle, product, n, box, fmg):
    """config 1 (tests/test_uniform_grid.f90), GSRB, Dirichlet planes from the exact solution."""
    orc, i_sol = P.uniform_problem(oracle.OracleMg, n, box)
    mg, _ = P.uniform_problem(product.Mg, n, box)
    h_o = P.run_cycles(orc, 5, fmg)
    h_g = P.run_cycles(mg, 5, fmg)
    check_histories(h_o, h_g)
    check_tolerance(orc, mg)
    assert_state_equal(orc, mg)
    # the reference test's own pass criterion: the error against the manufactured solution collapses
    assert P.max_error(mg, i_sol) < (1e-6 if fmg else 1e-4)
    if box == 16:
        assert_fast_path(mg)
    else:
        assert mg.stat("fast_launches") == 0 and mg.stat("general_launches") > 0     # box 8: general path only


@pytest.mark.parametrize("fmg", [False, True])
@pytest.mark.parametrize("smoother", [P.mg_smoother_gsrb, P.mg_smoother_gs])
def test_baseline_config0_uniform_128(oracle, product, smoother, fmg):
    """BASELINE.json configs[0] literally: tests/test_uniform_grid 128^3, box size 16, Dirichlet callback (the reference
    program's boundary condition), 10 iterations, FMG (`t`) and V-cycles (`f`), with the north-star's GSRB and
    with the program's own mg_smoother_gs.  588 boxes, levels 1..-5, four of them on the TMA-staged kernels."""
    orc, i_sol = P.uniform_problem(oracle.OracleMg, 128, 16, smoother=smoother)
    mg, _ = P.uniform_problem(product.Mg, 128, 16, smoother=smoother)
    assert mg.n_boxes == 588 and mg.lowest_lvl == -5
    h_o = P.run_cycles(orc, 10, fmg)
    h_g = P.run_cycles(mg, 10, fmg)
    check_histories(h_o, h_g)
    check_tolerance(orc, mg)
    assert_state_equal(orc, mg)
    # "max solution error" of the program after 10 iterations: FMG reaches rounding level, V-cycles ~0.1 per cycle
    assert P.max_error(mg, i_sol) < (1e-9 if fmg else 1e-6)
    if smoother == P.mg_smoother_gsrb:
        assert_fast_path(mg)
    else:                                      # the lexicographic in-box sweep runs on the general path; transfers stay fast
        assert_fast_path(mg, general_ok=True)


def test_baseline_config1_performance_256(oracle, product):
    """BASELINE.json configs[1] (tests/test_performance: rhs = 1, homogeneous Dirichlet, V(2,2) GSRB cycles) at 256^3 /
    box 16: 4684 boxes, five levels on the TMA-staged kernels -- the same program the bench times at 512^3, bitwise
    against the oracle incl. the full state (phi, rhs, old, res and phi's ghost faces) of every level."""
    orc = P.performance_problem(oracle.OracleMg, 256, 16)
    mg = P.performance_problem(product.Mg, 256, 16)
    assert mg.n_boxes == 4684      # 4096 + 512 + 64 + 8 + 1 boxes of 16^3, then 8^3, 4^3, 2^3
    check_histories(P.run_cycles(orc, 4, False), P.run_cycles(mg, 4, False))
    assert_state_equal(orc, mg)
    assert_fast_path(mg)


def test_performance_problem(oracle, product):
    """config 2 semantics at a size the oracle finishes quickly: rhs = 1, homogeneous Dirichlet."""
    orc = P.performance_problem(oracle.OracleMg, 64, 16)
    mg = P.performance_problem(product.Mg, 64, 16)
    check_histories(P.run_cycles(orc, 4, False), P.run_cycles(mg, 4, False))
    assert_state_equal(orc, mg)
    assert_fast_path(mg)


@pytest.mark.parametrize("op,bc,lam", [
    (P.mg_helmholtz, "neumann0", 1.0),
    (P.mg_helmholtz, "callback", 2.5),
    (P.mg_vlaplacian, "mixed", 0.0),
    (P.mg_vlaplacian, "callback", 0.0),
    (P.mg_vhelmholtz, "neumann0", 1.0),
    (P.mg_ahelmholtz, "mixed", 0.5),
    (P.mg_laplacian, "mixed", 0.0),
    (P.mg_laplacian, "continuous_x", 0.0),
])
@pytest.mark.parametrize("fmg", [False, True])
def test_operators_and_boundary_conditions(oracle, product, op, bc, lam, fmg):
    """config 4 semantics (variable-coefficient / Helmholtz operators, Neumann / mixed / continuous BC)."""
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, 32, 8, operator_type=op, bc=bc, lam=lam)
    check_histories(P.run_cycles(orc, 3, fmg), P.run_cycles(mg, 3, fmg))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("n,box", [((64, 32, 32), 8), ((48, 48, 48), 8), (32, 4), (64, 32), ((32, 64, 96), 16)])
def test_shapes_and_box_sizes(oracle, product, n, box):
    """non-cubic domains, several coarse boxes, box sizes 4 / 32, odd box counts."""
    orc, _ = P.uniform_problem(oracle.OracleMg, n, box)
    mg, _ = P.uniform_problem(product.Mg, n, box)
    check_histories(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("fmg", [False, True])
def test_refinement(oracle, product, fmg):
    """config 3 (tests/test_refinement.f90): octree AMR, refinement-boundary ghost interpolation."""
    orc, i_sol = P.refinement_problem(oracle.OracleMg, 3, 8, 32)
    mg, _ = P.refinement_problem(product.Mg, 3, 8, 32)
    # inputs built through the library's own restrict / ghost fill / operator must already agree
    assert_state_equal(orc, mg, variables=(2, i_sol), ghosts_of=(i_sol,))
    check_histories(P.run_cycles(orc, 4, fmg), P.run_cycles(mg, 4, fmg))
    assert_state_equal(orc, mg)
    assert P.max_error(mg, i_sol, leaves_only=True) < 1e-3


def test_refinement_nonzero_bc(oracle, product):
    """AMR with a non-zero Dirichlet value: exercises the zeroed-ghost state after mg_fas_fmg(have_guess=F)."""
    def build(cls):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb)
        for nb in range(1, 7):
            mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.75)
        mg.mg_set_methods()
        # refine a corner region so that refinement boundaries touch the physical boundary
        P.build_amr_tree(mg, 3, [32] * 3, 8, [1.0 / 32] * 3,
                         refine=lambda lvl, c, L: bool(np.all(c <= 0.5 ** lvl * L)))
        for lvl, id_ in P.all_ids(mg):
            mg.cc(id_)[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = 1.0 + 0.1 * lvl
        return mg
    orc, mg = build(oracle.OracleMg), build(product.Mg)
    check_histories(P.run_cycles(orc, 3, True), P.run_cycles(mg, 3, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)


def test_periodic_subtract_mean(oracle, product):
    """fully periodic Laplacian: subtract_mean is active (src/m_multigrid.f90:245-294).  The mean is a
    global sum whose association differs between a sequential CPU loop and a parallel reduction, so this
    case is held to the north-star tolerances instead of bitwise equality."""
    def build(cls):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb)
        mg.mg_set_methods()
        mg.mg_build_rectangle([32] * 3, 8, [1.0 / 32] * 3, [0.0] * 3, [True] * 3)
        mg.mg_load_balance()
        mg.mg_allocate_storage()
        for lvl, id_ in P.all_ids(mg):
            X, Y, Z = P.cell_coords(mg, id_, lvl)
            mg.cc(id_)[P.mg_irhs - 1] = np.sin(2 * np.pi * X) * np.cos(4 * np.pi * Y) + 0.3 * np.sin(2 * np.pi * Z) + 0.05
        return mg
    orc, mg = build(oracle.OracleMg), build(product.Mg)
    assert orc.get_int("subtract_mean") == 1 and mg.get_int("subtract_mean") == 1
    rhs_norm = max(float(np.abs(orc.cc(i)[P.mg_irhs - 1]).max()) for i in orc.ids(orc.highest_lvl))
    for fmg in (True, False):
        h_o, h_g = P.run_cycles(orc, 3, fmg), P.run_cycles(mg, 3, fmg)
        for a, b in zip(h_o, h_g):
            # a mean that differs in the last bit moves every residual by ~eps * |phi| / h^2: below that floor
            # the relative test is meaningless (SURVEY.md 8d "parity procedure")
            assert abs(a - b) <= max(RES_RTOL * abs(a), 1e-12 * rhs_norm)
    d = state_diff(orc, mg, variables=(1,), ghosts_of=())["phi"]
    assert d[0] <= 1e-12 * d[1]


def test_phi_bc_store(oracle, product):
    """mg_phi_bc_store (src/m_ghost_cells.f90:66-117): planes kept in the rhs ghost cells."""
    orc, _ = P.uniform_problem(oracle.OracleMg, 32, 8)
    mg, _ = P.uniform_problem(product.Mg, 32, 8)
    orc.mg_phi_bc_store()
    mg.mg_phi_bc_store()
    check_histories(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    assert_state_equal(orc, mg)


def test_resident_mode_and_graph_replay(oracle, product):
    """device-resident cycles (omg_fas_vcycles, CUDA-graph replay) give the same numbers as call-by-call."""
    orc = P.performance_problem(oracle.OracleMg, 64, 16)
    mg = P.performance_problem(product.Mg, 64, 16, gpu_resident=True)
    mg.mg_gpu_push(P.mg_iphi)
    mg.mg_gpu_push(P.mg_irhs)
    mg.mg_fas_vcycle(want_res=False)          # uploads parameters / boundary data
    h_o = P.run_cycles(orc, 5, False)
    h_g, ms = mg.fas_vcycles(4)
    assert h_g == h_o[1:]
    assert mg.stat("graph_launches") >= 4
    for iv in (1, 2, 3, 4):
        mg.mg_gpu_pull(iv)
    assert_state_equal(orc, mg)


def test_no_cpu_fallback_errors(product):
    """error behaviour of the boundary (the reference error-stops; the C ABI returns codes + messages)."""
    mg = product.Mg(P.mg_laplacian, P.mg_smoother_gsrb)
    mg.mg_set_methods()
    mg.mg_build_rectangle([32] * 3, 8, [1.0 / 32] * 3, [0.0] * 3, [False] * 3)
    mg.mg_load_balance()
    mg.mg_allocate_storage()
    mg.set_int("smoother_type", 7)
    with pytest.raises(product.MgError, match="unsupported smoother"):
        mg.mg_fas_vcycle()
    mg.set_int("smoother_type", P.mg_smoother_gsrb)
    with pytest.raises(product.MgError, match="lvl > highest_lvl"):
        mg.mg_fill_ghost_cells_lvl(5, P.mg_iphi)


@pytest.mark.parametrize("fmg", [False, True])
@pytest.mark.parametrize("op,bc,lam,n,box", [
    (P.mg_laplacian, "callback", 0.0, 32, 8),       # the literal setting of tests/test_uniform_grid.f90:64
    (P.mg_laplacian, "callback", 0.0, 64, 16),
    (P.mg_helmholtz, "neumann0", 1.0, 32, 8),
    (P.mg_vlaplacian, "mixed", 0.0, 32, 8),
    (P.mg_ahelmholtz, "mixed", 0.5, 32, 4),
])
def test_gs_smoother(oracle, product, op, bc, lam, n, box, fmg):
    """mg_smoother_gs, the reference's default and what its test programs hard-code: lexicographic
    Gauss-Seidel inside each box against ghost cells filled before the sweep.  The device sweeps hyperplanes
    i + j + k = const, which performs the same arithmetic on the same values: bitwise equal to the oracle."""
    orc, i_sol = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, bc=bc, lam=lam, smoother=P.mg_smoother_gs)
    mg, _ = P.uniform_problem(product.Mg, n, box, operator_type=op, bc=bc, lam=lam, smoother=P.mg_smoother_gs)
    assert mg.lowest_lvl == orc.lowest_lvl
    check_histories(P.run_cycles(orc, 3, fmg), P.run_cycles(mg, 3, fmg))
    assert_state_equal(orc, mg)


def test_gs_smoother_refinement(oracle, product):
    """tests/test_refinement.f90 as shipped (mg_smoother_gs, :65) on the AMR grid."""
    orc, i_sol = P.refinement_problem(oracle.OracleMg, 3, 8, 32, smoother=P.mg_smoother_gs)
    mg, _ = P.refinement_problem(product.Mg, 3, 8, 32, smoother=P.mg_smoother_gs)
    check_histories(P.run_cycles(orc, 3, True), P.run_cycles(mg, 3, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)


@pytest.mark.parametrize("op,bc,lam,n,box", [
    (P.mg_vlaplacian, "mixed", 0.0, 64, 16),
    (P.mg_vlaplacian, "callback", 0.0, (64, 32, 32), 16),
    (P.mg_vhelmholtz, "neumann0", 1.0, 64, 16),
    (P.mg_vhelmholtz, "mixed", 0.5, 64, 32),
    (P.mg_helmholtz, "neumann0", 1.0, 64, 16),
    (P.mg_ahelmholtz, "mixed", 0.75, 64, 16),               # three coefficient arrays (NE = 3 kernels, half-box tiles)
    (P.mg_ahelmholtz, "callback", 1.5, (64, 32, 32), 16),
    (P.mg_ahelmholtz, "neumann0", 1.0, 64, 32),
])
def test_fast_kernels_variable_coefficient(oracle, product, op, bc, lam, n, box):
    """config 4 at box sizes 16 / 32: the TMA-staged kernels of the variable-coefficient operators (eps tile + stored
    eps ghost faces, harmonic-mean face coefficients recomputed per cell) against the oracle, bitwise."""
    orc, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, n, box, operator_type=op, bc=bc, lam=lam)
    check_histories(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    assert_fast_path(mg, general_ok="small")


@pytest.mark.parametrize("op,bc,lam,n,box", [
    (P.mg_vlaplacian, "mixed", 0.0, 64, 16),
    (P.mg_vhelmholtz, "neumann0", 1.0, (64, 32, 32), 16),
    (P.mg_ahelmholtz, "mixed", 0.75, 64, 16),
    (P.mg_vlaplacian, "callback", 0.0, 64, 32),
])
def test_tolerance_mode_of_the_coefficient_kernels(oracle, product, op, bc, lam, n, box):
    """option coeff_math = 1: the TMA-staged coefficient kernels take their six face coefficients from one division by
    batch inversion (omg_device.cuh recip6).  NOT bitwise: held to the north-star tolerances instead -- per-cycle residuals
    within 1e-10 relative (plus the rounding floor of a converged residual), final solution within 1e-12 * |phi|"""
    orc, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, bc=bc, lam=lam)
    mg, _ = P.uniform_problem(product.Mg, n, box, operator_type=op, bc=bc, lam=lam)
    mg.set_option("coeff_math", 1)
    h_o = P.run_cycles(orc, 2, True) + P.run_cycles(orc, 3, False)
    h_g = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 3, False)
    dr = np.asarray(mg.dr(mg.highest_lvl))
    floor = 64 * np.finfo(float).eps * 2.0 * float(np.sum(1.0 / dr ** 2)) * 4.0       # eps up to ~2, |phi| ~ 1
    for a, b in zip(h_o, h_g):
        assert abs(a - b) <= 1e-10 * abs(a) + floor, (h_o, h_g, floor)
    diff = state_diff(orc, mg, variables=(1,), ghosts_of=())
    assert diff["phi"][0] <= 1e-12 * max(diff["phi"][1], 1e-300), diff
    assert_fast_path(mg, general_ok="small")


@pytest.mark.parametrize("ring", [0, 2])
@pytest.mark.parametrize("case", ["uniform_callback", "helmholtz_mixed", "noncubic_dirichlet0", "amr"])
def test_ring_smoother_options(oracle, product, case, ring):
    """option ring: 2 = the persistent ring smoother (k_smooth_ring: producer warp + 5-stage mbarrier ring) on EVERY box-16
    level it serves, 0 = one tile per CTA (k_smooth_lpl_fast).  Both against the oracle, bitwise, on stored boundary planes,
    Neumann / Dirichlet constants, a non-cubic domain (several boxes on the coarse levels) and refinement boundaries."""
    def build(cls):
        if case == "uniform_callback":
            return P.uniform_problem(cls, 64, 16)[0]
        if case == "helmholtz_mixed":
            return P.uniform_problem(cls, 64, 16, operator_type=P.mg_helmholtz, bc="mixed", lam=1.5)[0]
        if case == "noncubic_dirichlet0":
            return P.uniform_problem(cls, (96, 64, 32), 16, bc="dirichlet0")[0]
        return P.refinement_problem(cls, 3, 16, 64)[0]
    orc, mg = build(oracle.OracleMg), build(product.Mg)
    mg.set_option("ring", ring)
    check_histories(P.run_cycles(orc, 2, True), P.run_cycles(mg, 2, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    assert mg.stat("fast_launches") > 0


def test_store_fine_residual_off(oracle, product):
    """option store_fine_residual = 0 (device-resident hosts that never read mg_ires): the cycle's numbers do not change --
    residual histories, phi, rhs, old bitwise -- only the level's own residual array is no longer written"""
    orc = P.performance_problem(oracle.OracleMg, 64, 16)
    mg = P.performance_problem(product.Mg, 64, 16)
    mg.set_option("store_fine_residual", 0)
    check_histories(P.run_cycles(orc, 3, True), P.run_cycles(mg, 3, True))
    check_histories(P.run_cycles(orc, 3, False), P.run_cycles(mg, 3, False))
    assert_state_equal(orc, mg, variables=(1, 2, 3))
    d = state_diff(orc, mg, variables=(4,), ghosts_of=(), levels=[mg.highest_lvl])["res"]
    assert d[0] > 0.0      # the finest level's residual really was not stored


def test_refinement_on_fast_kernels(oracle, product):
    """BASELINE.json configs[2] with boxes of 16 (`test_refinement 3 16 64 64 64`): levels with refinement-boundary faces
    (FK_PLANE, filled by sides_rb) run the TMA-staged kernels too -- only k_rb_fill / k_oldpl are general-path helpers"""
    orc, i_sol = P.refinement_problem(oracle.OracleMg, 3, 16, 64)
    mg, _ = P.refinement_problem(product.Mg, 3, 16, 64)
    f0, g0 = mg.stat("fast_launches"), mg.stat("general_launches")
    check_histories(P.run_cycles(orc, 3, True), P.run_cycles(mg, 3, True))
    check_histories(P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert_state_equal(orc, mg)
    assert P.max_error(mg, i_sol, leaves_only=True) < 1e-3
    assert mg.stat("fast_launches") > f0 and mg.stat("general_launches") == g0


def test_golden_fixtures_on_the_device(product):
    """the committed fixtures (tests/golden/golden.json, oracle output): residual histories as exact hex floats and a
    position-weighted checksum of phi on the leaves, reproduced by the CUDA path without running the oracle"""
    import json
    import os

    import golden.make_golden as mk
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")) as f:
        gold = json.load(f)
    assert len(gold["cases"]) >= 10
    for case in gold["cases"]:
        got = mk.run_case(product.Mg, case["name"])
        assert got["residuals"] == case["residuals"], case["name"]
        assert got["phi_checksum"] == case["phi_checksum"], case["name"]
