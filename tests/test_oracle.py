"""CPU tests of the oracle itself.  The reference ships no golden vectors and cannot be compiled here
(SURVEY.md F2, F3), so the oracle is pinned by (i) the reference tests' own pass criterion and the
surveyor's convergence anchors, (ii) rank-count invariance, (iii) GSRB box-size invariance, (iv) bitwise
agreement with an independent global-array restatement, (v) committed golden fixtures (regression)."""
import json
import os

import numpy as np
import pytest

import problems as P
from compare import assert_state_equal

HERE = os.path.dirname(os.path.abspath(__file__))


def global_phi(mg, lvl=None, iv=P.mg_iphi):
    """assemble the finest uniform level into one [z, y, x] array"""
    lvl = mg.highest_lvl if lvl is None else lvl
    nc = mg.box_size_lvl(lvl)
    ids = mg.ids(lvl)
    ix = np.array([mg.box_info(i)["ix"] for i in ids])
    nb = ix.max(axis=0)
    out = np.zeros((nb[2] * nc, nb[1] * nc, nb[0] * nc))
    for id_, (bx, by, bz) in zip(ids, ix):
        out[(bz - 1) * nc:bz * nc, (by - 1) * nc:by * nc, (bx - 1) * nc:bx * nc] = mg.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
    return out


# max error per iteration, 64^3, GSRB, Dirichlet callback
ANCHOR_V = [9.5e-2, 8.1e-3, 6.8e-4, 5.8e-5, 4.9e-6, 4.2e-7]
ANCHOR_FMG = [2.6e-2, 4.6e-4, 8.8e-6, 2.2e-7, 4.8e-9, 9.8e-11]


@pytest.mark.parametrize("fmg,anchor", [(False, ANCHOR_V), (True, ANCHOR_FMG)])
def test_convergence_anchors(oracle, fmg, anchor):
    mg, i_sol = P.uniform_problem(oracle.OracleMg, 64, 16)
    for it, ref in enumerate(anchor, 1):
        mg.mg_fas_fmg(it > 1) if fmg else mg.mg_fas_vcycle()
        err = P.max_error(mg, i_sol)
        assert abs(err - ref) <= 0.03 * ref, (it, err, ref)


def test_manufactured_solution_reaches_rounding(oracle):
    """tests/test_uniform_grid.f90's criterion: rhs is the discrete operator of the exact field, so the
    error must fall to rounding level, not O(h^2)."""
    mg, i_sol = P.uniform_problem(oracle.OracleMg, 32, 8)
    hist = P.run_cycles(mg, 9, True)
    assert P.max_error(mg, i_sol) < 5e-13
    assert hist[-1] < 1e-9 * hist[0]


@pytest.mark.parametrize("n_cpu", [2, 3, 8])
def test_rank_count_invariance_uniform(oracle, n_cpu):
    """results are bitwise independent of the number of MPI ranks."""
    a, _ = P.uniform_problem(oracle.OracleMg, 32, 8)
    b, _ = P.uniform_problem(oracle.OracleMg, 32, 8, n_cpu=n_cpu)
    assert P.run_cycles(a, 2, True) == P.run_cycles(b, 2, True)
    assert P.run_cycles(a, 2, False) == P.run_cycles(b, 2, False)
    assert_state_equal(a, b)
    ranks = {a_["rank"] for a_ in (b.box_info(i) for i in b.ids(1))}
    assert len(ranks) == n_cpu


@pytest.mark.parametrize("n_cpu", [2, 5])
def test_rank_count_invariance_amr(oracle, n_cpu):
    a, _ = P.refinement_problem(oracle.OracleMg, 3, 8, 32)
    b, _ = P.refinement_problem(oracle.OracleMg, 3, 8, 32, n_cpu=n_cpu)
    assert P.run_cycles(a, 2, True) == P.run_cycles(b, 2, True)
    assert P.run_cycles(a, 2, False) == P.run_cycles(b, 2, False)
    assert_state_equal(a, b)


def test_gsrb_box_size_invariance(oracle):
    """with GSRB on a uniform grid the result does not depend on the box size."""
    ref = None
    for box in (4, 8, 16, 32):
        mg, _ = P.uniform_problem(oracle.OracleMg, 32, box)
        hist = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
        g = global_phi(mg)
        if ref is None:
            ref = (hist, g)
        else:
            assert hist == ref[0]
            assert np.array_equal(g, ref[1])


@pytest.mark.parametrize("helm", [False, True])
def test_global_array_restatement_agrees_bitwise(oracle, helm):
    """independent numpy multigrid on one global array per level == box-structured oracle, bit for bit"""
    from oracle.global_array_mg import GlobalMG
    n, box, lam = 32, 8, (1.5 if helm else 0.0)
    op = P.mg_helmholtz if helm else P.mg_laplacian
    mg, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, lam=lam)

    def bc_fn(l, nb, n_l, dr_l):
        return P.mg_bc_dirichlet, P.face_plane(P.solution, np.zeros(3), np.full(3, dr_l), n_l, nb)
    G = GlobalMG(n, 1.0 / n, bc_fn, lam=lam, helmholtz=helm)
    G.rhs[0][1:-1, 1:-1, 1:-1] = global_phi(mg, iv=P.mg_irhs)
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    assert h_box == h_glob
    assert np.array_equal(global_phi(mg), G.phi[0][1:-1, 1:-1, 1:-1])


@pytest.mark.parametrize("op,lam,bc", [(P.mg_vlaplacian, 0.0, "callback"), (P.mg_vhelmholtz, 0.75, "mixed"), (P.mg_ahelmholtz, 1.25, "mixed")])
def test_global_array_restatement_variable_coefficients(oracle, op, lam, bc):
    """the same for vlaplacian / vhelmholtz / ahelmholtz: harmonic-mean face coefficients in the reference's expression
    order on one global array per level (coefficient arrays, ghost layer included, from the same analytic fields), bit
    for bit against the box-structured oracle (V-cycles and FMG)"""
    from oracle.global_array_mg import GlobalMG
    n, box = 32, 8
    mg, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, lam=lam, bc=bc)

    def bc_fn(l, nb, n_l, dr_l):
        if bc == "callback":
            return P.mg_bc_dirichlet, P.face_plane(P.solution, np.zeros(3), np.full(3, dr_l), n_l, nb)
        return (P.mg_bc_dirichlet, 0.25) if nb == 1 else (P.mg_bc_neumann, 0.5)

    def eps_fn(l, n_l, dr_l):
        x = (np.arange(n_l + 2) - 0.5) * dr_l           # ghost layer included; exact in binary for these sizes
        Z, Y, X = np.meshgrid(x, x, x, indexing="ij")
        return [P.eps_field(X, Y, Z, d) for d in range(P.n_eps(op))]
    G = GlobalMG(n, 1.0 / n, bc_fn, lam=lam, helmholtz=(op != P.mg_vlaplacian), eps_fn=eps_fn)
    # the oracle's coefficient arrays are these very fields (checked on the finest level, interior and ghost layer)
    for id_ in mg.ids(mg.highest_lvl):
        X, Y, Z = P.cell_coords(mg, id_, mg.highest_lvl)
        for d in range(P.n_eps(op)):
            assert np.array_equal(mg.cc(id_)[P.mg_iveps - 1 + d], P.eps_field(X, Y, Z, d))
    G.rhs[0][1:-1, 1:-1, 1:-1] = global_phi(mg, iv=P.mg_irhs)
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    assert h_box == h_glob
    assert np.array_equal(global_phi(mg), G.phi[0][1:-1, 1:-1, 1:-1])



@pytest.mark.parametrize("helm,box,bc", [(False, 8, "callback"), (True, 16, "mixed"), (False, 4, "mixed")])
def test_global_array_restatement_gs_smoother(oracle, helm, box, bc):
    """mg_smoother_gs, the reference's default: the result depends on the box size (a box reads its neighbours' values of
    before the sweep), so the global-array restatement takes the box size and sweeps each box with plain nested loops:
    bit for bit against the box-structured oracle, V-cycles and FMG"""
    from oracle.global_array_mg import GlobalMG
    n, lam = 32, (1.5 if helm else 0.0)
    op = P.mg_helmholtz if helm else P.mg_laplacian
    mg, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=op, lam=lam, bc=bc, smoother=P.mg_smoother_gs)

    def bc_fn(l, nb, n_l, dr_l):
        if bc == "callback":
            return P.mg_bc_dirichlet, P.face_plane(P.solution, np.zeros(3), np.full(3, dr_l), n_l, nb)
        return (P.mg_bc_dirichlet, 0.25) if nb == 1 else (P.mg_bc_neumann, 0.5)
    G = GlobalMG(n, 1.0 / n, bc_fn, lam=lam, helmholtz=helm, smoother="gs", box=box)
    G.rhs[0][1:-1, 1:-1, 1:-1] = global_phi(mg, iv=P.mg_irhs)
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    assert h_box == h_glob
    assert np.array_equal(global_phi(mg), G.phi[0][1:-1, 1:-1, 1:-1])
    # and it does depend on the box size: another decomposition gives other bits
    G2 = GlobalMG(n, 1.0 / n, bc_fn, lam=lam, helmholtz=helm, smoother="gs", box=(8 if box != 8 else 16))
    G2.rhs[0][...] = G.rhs[0]
    assert G2.fmg(False) != h_glob[0]


@pytest.mark.parametrize("smoother", ["gsrb", "gs"])
def test_global_array_restatement_periodic_subtract_mean(oracle, smoother):
    """fully periodic Laplacian with subtract_mean (src/m_multigrid.f90:245-294) on the global arrays.  The mean is a sum
    whose association differs (box by box in the oracle, one numpy sum here): north-star tolerances instead of bits"""
    from oracle.global_array_mg import GlobalMG
    n, box = 32, 8
    mg = oracle.OracleMg(P.mg_laplacian, P.mg_smoother_gs if smoother == "gs" else P.mg_smoother_gsrb)
    mg.mg_set_methods()
    mg.mg_build_rectangle([n] * 3, box, [1.0 / n] * 3, [0.0] * 3, [True] * 3)
    mg.mg_load_balance()
    mg.mg_allocate_storage()
    assert mg.get_int("subtract_mean") == 1
    for lvl, id_ in P.all_ids(mg):
        X, Y, Z = P.cell_coords(mg, id_, lvl)
        mg.cc(id_)[P.mg_irhs - 1] = np.sin(2 * np.pi * X) * np.cos(4 * np.pi * Y) + 0.3 * np.sin(2 * np.pi * Z) + 0.05
    G = GlobalMG(n, 1.0 / n, None, smoother=smoother, box=box, periodic=True)
    G.rhs[0][1:-1, 1:-1, 1:-1] = global_phi(mg, iv=P.mg_irhs)
    rhs_norm = float(np.abs(G.rhs[0]).max())
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    for a, b in zip(h_box, h_glob):
        assert abs(a - b) <= max(1e-10 * abs(a), 1e-12 * rhs_norm), (h_box, h_glob)
    assert h_box[-1] < 1e-3 * h_box[0]
    a, b = global_phi(mg), G.phi[0][1:-1, 1:-1, 1:-1]
    assert np.abs(a - b).max() <= 1e-12 * np.abs(a).max()
    assert abs(a.mean()) < 1e-15 and abs(b.mean()) < 1e-15


def test_global_array_neumann_constants(oracle):
    from oracle.global_array_mg import GlobalMG
    n, box = 32, 16
    mg, _ = P.uniform_problem(oracle.OracleMg, n, box, operator_type=P.mg_helmholtz, lam=1.0, bc="mixed")
    G = GlobalMG(n, 1.0 / n, lambda l, nb, n_l, dr_l: (P.mg_bc_dirichlet, 0.25) if nb == 1 else (P.mg_bc_neumann, 0.5),
                 lam=1.0, helmholtz=True)
    G.rhs[0][1:-1, 1:-1, 1:-1] = global_phi(mg, iv=P.mg_irhs)
    assert P.run_cycles(mg, 3, True) == [G.fmg(False), G.fmg(True), G.fmg(True)]
    assert np.array_equal(global_phi(mg), G.phi[0][1:-1, 1:-1, 1:-1])


@pytest.mark.parametrize("op,bc,lam", [(P.mg_vlaplacian, "mixed", 0.0), (P.mg_vhelmholtz, "neumann0", 1.0),
                                       (P.mg_ahelmholtz, "mixed", 0.5), (P.mg_helmholtz, "neumann0", 1.0)])
def test_operators_converge(oracle, op, bc, lam):
    mg, _ = P.uniform_problem(oracle.OracleMg, 32, 8, operator_type=op, bc=bc, lam=lam)
    hist = P.run_cycles(mg, 6, True)
    assert hist[-1] < 1e-4 * hist[0]
    assert all(b < a for a, b in zip(hist, hist[1:]))


def test_gs_smoother_converges(oracle):
    """the reference default mg_smoother_gs (used by its test programs)"""
    mg, i_sol = P.uniform_problem(oracle.OracleMg, 32, 8, smoother=P.mg_smoother_gs)
    P.run_cycles(mg, 6, True)
    assert P.max_error(mg, i_sol) < 1e-6


def test_refinement_converges(oracle):
    mg, i_sol = P.refinement_problem(oracle.OracleMg, 3, 8, 32)
    assert [len(mg.ids(l)) for l in (1, 2, 3)] == [64, 64, 64]
    e0 = P.max_error(mg, i_sol, leaves_only=True)
    P.run_cycles(mg, 8, True)
    assert P.max_error(mg, i_sol, leaves_only=True) < 1e-10 * e0


def test_level_structure_matches_survey_appendix_b(oracle):
    """SURVEY.md appendix B (restated from src/m_build_tree.f90:51-72)"""
    expect = {(64, 16): (-4, -1, 76), (128, 16): (-5, -2, 588), (128, 32): (-5, -1, 77)}
    for (n, box), (lo, fn, nb) in expect.items():
        mg = oracle.OracleMg()
        mg.mg_set_methods()
        mg.mg_build_rectangle([n] * 3, box, [1.0 / n] * 3, [0.0] * 3, [False] * 3)
        assert (mg.lowest_lvl, mg.first_normal_lvl, mg.n_boxes) == (lo, fn, nb)


def test_error_conditions(oracle):
    mg = oracle.OracleMg()
    with pytest.raises(oracle.OracleError, match="box_size should be even"):
        mg.mg_build_rectangle([30] * 3, 5, [1.0] * 3, [0.0] * 3, [False] * 3)
    with pytest.raises(oracle.OracleError, match="does not divide"):
        mg.mg_build_rectangle([30] * 3, 8, [1.0] * 3, [0.0] * 3, [False] * 3)
    mg2, _ = P.uniform_problem(oracle.OracleMg, 16, 8)
    with pytest.raises(oracle.OracleError, match="lvl > highest_lvl"):
        mg2.mg_fill_ghost_cells_lvl(3, P.mg_iphi)
    with pytest.raises(oracle.OracleError, match="cannot prolong highest"):
        mg2.mg_prolong(1, P.mg_iphi, P.mg_iphi, False)


def test_golden_fixtures(oracle):
    """regression fixtures generated by tests/golden/make_golden.py from the oracle"""
    path = os.path.join(HERE, "golden", "golden.json")
    with open(path) as f:
        gold = json.load(f)
    import golden.make_golden as mk
    for case in gold["cases"]:
        got = mk.run_case(oracle.OracleMg, case["name"])
        assert got["residuals"] == case["residuals"], case["name"]
        assert got["phi_checksum"] == case["phi_checksum"], case["name"]


def _level_arrays(mg, lvl, iv):
    """(array over the whole domain at the resolution of `lvl` filled where boxes exist, lo, hi of the boxes' bounding patch)"""
    nc = mg.box_size_lvl(lvl)
    ids = mg.ids(lvl)
    ix = np.array([mg.box_info(i)["ix"] for i in ids])
    n = [int(round(e / mg.dr(lvl)[0])) for e in getattr(mg, "_domain_extent", (1.0, 1.0, 1.0))]
    out = np.zeros((n[2], n[1], n[0]))
    for id_, (bx, by, bz) in zip(ids, ix):
        out[(bz - 1) * nc:bz * nc, (by - 1) * nc:by * nc, (bx - 1) * nc:bx * nc] = mg.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
    lo, hi = (ix.min(axis=0) - 1) * nc, ix.max(axis=0) * nc
    assert len(ids) * nc ** 3 == int(np.prod(hi - lo)), "the level is not one rectangular patch"
    return out, lo, hi


AMR_CASES = {
    # the reference's own rule (tests/test_refinement.f90:191-247): centred cubes, three levels
    "centred": (3, 8, 32, None, [(P.mg_bc_dirichlet, 0.0)] * 6),
    # a patch that touches the domain boundary on five sides (the coarse-side interpolation reads physical ghost cells),
    # Neumann / non-zero Dirichlet constants, Helmholtz, and a second refinement inside it
    "slab": (3, 4, 16, lambda lvl, c, L: bool(c[0] < (0.5 if lvl == 1 else 0.25) * L[0]),
             [(P.mg_bc_dirichlet, 0.25), (P.mg_bc_neumann, 0.5), (P.mg_bc_neumann, -0.25), (P.mg_bc_dirichlet, -1.0),
              (P.mg_bc_continuous, 0.0), (P.mg_bc_dirichlet, 0.0)]),
}


@pytest.mark.parametrize("case", sorted(AMR_CASES))
@pytest.mark.parametrize("helm", [False, True])
def test_global_array_restatement_refined_grid(oracle, case, helm):
    """independent numpy multigrid with one global array per level and rectangular refined patches
    (oracle/global_array_amr.py) == box-structured oracle on the same refined tree, bit for bit: FMG and V-cycle residual
    histories and phi on every level.  Pins the refinement-boundary ghost cells (box_gc_for_fine_neighbor + sides_rb), the
    partial restriction / coarse right-hand side / correction and the parent / leaf bookkeeping."""
    from oracle.global_array_amr import GlobalAMR
    n_levels, box, n, refine, bc = AMR_CASES[case]
    lam = 1.5 if helm else 0.0
    mg = oracle.OracleMg(P.mg_helmholtz if helm else P.mg_laplacian, P.mg_smoother_gsrb)
    if helm:
        mg.set_lambda(lam)
    for nb in range(1, 7):
        mg.set_bc(nb, P.mg_iphi, *bc[nb - 1])
    mg.mg_set_methods()
    P.build_amr_tree(mg, n_levels, [n] * 3, box, [1.0 / n] * 3, refine=refine)
    rng = np.random.default_rng(7)
    levels = list(range(mg.lowest_lvl, mg.highest_lvl + 1))
    for lvl in levels:
        for id_ in mg.ids(lvl):
            mg.cc(id_)[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((mg.box_size_lvl(lvl),) * 3)
    patches = []
    for lvl in range(2, mg.highest_lvl + 1):
        _, lo, hi = _level_arrays(mg, lvl, P.mg_irhs)
        patches.append((lo, hi))
    G = GlobalAMR(n, 1.0 / n, patches, bc, lam=lam, helmholtz=helm)
    assert G.nl == len(levels)
    for l, lvl in enumerate(levels):
        a, lo, hi = _level_arrays(mg, lvl, P.mg_irhs)
        assert np.array_equal(lo, G.lo[l]) and np.array_equal(hi, G.hi[l])
        G.rhs[l][1:-1, 1:-1, 1:-1] = a
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    assert h_box == h_glob
    for l, lvl in enumerate(levels):
        a, lo, hi = _level_arrays(mg, lvl, P.mg_iphi)
        s = tuple(slice(lo[d], hi[d]) for d in (2, 1, 0))
        assert np.array_equal(a[s], G.phi[l][1:-1, 1:-1, 1:-1][s]), lvl


@pytest.mark.parametrize("dims,box,bc_name,helm", [((32, 16, 16), 8, "mixed", False), ((32, 16, 16), 8, "mixed", True),
                                                   ((16, 32, 16), 4, "neumann0", True),      # (all-Neumann Laplacian: singular)
                                                   ((32, 32, 16), 8, "dirichlet0", False), ((32, 32, 16), 8, "dirichlet0", True)])
def test_global_array_restatement_non_cubic_grid(oracle, dims, box, bc_name, helm):
    """non-cubic domains: the coarsening stops when one direction reaches two cells, so the coarsest level -- the one the
    coarse solve iterates on (src/m_multigrid.f90:202-208) -- has 2 x 1 x 1 or 2 x 2 x 1 boxes.  The global-array multigrid
    (its base level a brick) against the box-structured oracle, bit for bit: residual histories of FMG and V-cycles, phi
    on every level"""
    from oracle.global_array_amr import GlobalAMR
    lam = 1.5 if helm else 0.0
    mg, _ = P.uniform_problem(oracle.OracleMg, dims, box, operator_type=P.mg_helmholtz if helm else P.mg_laplacian, bc=bc_name, lam=lam)
    mg._domain_extent = tuple(d / dims[0] for d in dims)
    bc = {"mixed": [(P.mg_bc_dirichlet, 0.25)] + [(P.mg_bc_neumann, 0.5)] * 5, "neumann0": [(P.mg_bc_neumann, 0.0)] * 6,
          "dirichlet0": [(P.mg_bc_dirichlet, 0.0)] * 6}[bc_name]
    G = GlobalAMR(dims, 1.0 / dims[0], [], bc, lam=lam, helmholtz=helm)
    levels = list(range(mg.lowest_lvl, mg.highest_lvl + 1))
    assert G.nl == len(levels) and min(G.ns[0]) == 2 and max(G.ns[0]) > 2
    a, lo, hi = _level_arrays(mg, mg.highest_lvl, P.mg_irhs)
    G.rhs[-1][1:-1, 1:-1, 1:-1] = a
    h_box = P.run_cycles(mg, 2, True) + P.run_cycles(mg, 2, False)
    h_glob = [G.fmg(False), G.fmg(True), G.vcycle(), G.vcycle()]
    assert h_box == h_glob
    for l, lvl in enumerate(levels):
        a, lo, hi = _level_arrays(mg, lvl, P.mg_iphi)
        assert np.array_equal(a, G.phi[l][1:-1, 1:-1, 1:-1]), lvl
