"""On-device regrid / re-partition: omg_update_tree behind mg_gpu_update_tree.

The reference changes a tree with mg_add_children + mg_set_neighbors_lvl + mg_set_refinement_boundaries
(src/m_build_tree.f90:296-367) and mg_load_balance (src/m_load_balance.f90:71-136); an application then deallocates,
rebuilds and copies every variable in again (coupling_amrvac).  Here the tree of an ALLOCATED mg is refined by one more
level and re-balanced; boxes that exist before and after must keep their device data (copied locally, or sent from the
old to the new owner), new boxes start from zero.  The oracle takes the reference's route: a fresh tree built by the same
sequence of calls (so box ids agree), every variable of the surviving boxes copied over on the host.  After that both
continue to cycle; everything is compared bitwise."""
import functools

import numpy as np
import pytest

import problems as P
from compare import assert_state_equal, assert_trees_equal

pytestmark = pytest.mark.gpu

N, BOX = 32, 8


def in_region(lvl, center, domain_len):
    r0 = 0.5 * domain_len - domain_len * 0.5 ** (lvl + 1)
    r1 = 0.5 * domain_len + domain_len * 0.5 ** (lvl + 1)
    return bool(np.all((center >= r0) & (center <= r1)))


def refine_level(mg, lvl):
    """children for the boxes of `lvl` inside the centred region, then the bookkeeping of m_build_tree for lvl + 1"""
    domain_len = np.array([1.0, 1.0, 1.0])
    for id_ in mg.ids(lvl):
        center = mg.box_info(id_)["r_min"] + 0.5 * BOX * mg.dr(lvl)
        if in_region(lvl, center, domain_len):
            mg.mg_add_children(id_)
    mg.mg_set_leaves_parents(lvl)
    mg.mg_set_next_level_ids(lvl)
    mg.mg_set_neighbors_lvl(lvl + 1)


def finish_tree(mg, highest, lb="default"):
    mg.mg_set_leaves_parents(highest)
    mg.highest_lvl = highest
    for lvl in range(1, highest + 1):
        mg.mg_set_refinement_boundaries(lvl)
    if lb == "simple":      # two of the reference's partitioners: switching between them moves existing boxes
        mg.mg_load_balance_simple()
    else:
        mg.mg_load_balance()
        mg.mg_load_balance_parents()


def new_mg(cls, **kw):
    mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, n_extra_vars=1, **kw)
    for nb in range(1, 7):
        mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.125 * nb)
        mg.set_bc(nb, P.mg_num_vars + 1, P.mg_bc_neumann, 0.0)
    mg.mg_set_methods()
    mg.mg_build_rectangle([N] * 3, BOX, [1.0 / N] * 3, [0.0] * 3, [False] * 3, 4000)
    return mg


def set_rhs(mg, ids=None):
    for lvl, id_ in (P.all_ids(mg) if ids is None else ids):
        X, Y, Z = P.cell_coords(mg, id_, lvl)
        cc = mg.cc(id_)
        cc[P.mg_irhs - 1] = np.sin(2.2 * X + 0.3) * np.cos(3.1 * Y) + 0.4 * Z + 0.1 * lvl
        cc[P.mg_num_vars] = 1.0 + 0.5 * X * Y - 0.25 * Z      # an extra variable that must travel too


@pytest.mark.parametrize("n_cpu", [1, 2, 3])
def test_refine_and_rebalance_an_allocated_tree(oracle, product, n_cpu):
    kw = {"n_cpu": n_cpu} if n_cpu > 1 else {}
    # ---- stage A: two levels, allocated, a few cycles ----
    orc_a = new_mg(oracle.OracleMg, **kw)
    mg = new_mg(product.Mg if n_cpu == 1 else functools.partial(product.MgGroup, n_cpu))
    for m in (orc_a, mg):
        refine_level(m, 1)
        finish_tree(m, 2, lb="simple")
        m.mg_allocate_storage()
        set_rhs(m)
    h_a = (P.run_cycles(orc_a, 2, True), P.run_cycles(mg, 2, True))
    assert h_a[0] == h_a[1]
    n_old = orc_a.n_boxes
    old_ids = P.all_ids(orc_a)
    # ---- stage B, product: one more level on the allocated tree, new partition, on-device regrid ----
    refine_level(mg, 2)
    finish_tree(mg, 3)
    mg.mg_gpu_update_tree()
    # ---- stage B, oracle: the reference's route (fresh tree by the same calls, every variable copied in again) ----
    orc = new_mg(oracle.OracleMg, **kw)
    refine_level(orc, 1)
    refine_level(orc, 2)
    finish_tree(orc, 3)
    orc.mg_allocate_storage()
    assert_trees_equal(orc, mg if n_cpu == 1 else mg.ranks[0])
    assert orc.n_boxes > n_old
    for lvl, id_ in old_ids:
        orc.cc(id_)[...] = orc_a.cc(id_)
    moved = sum(1 for lvl, id_ in old_ids if orc_a.box_info(id_)["rank"] != orc.box_info(id_)["rank"])
    assert n_cpu != 3 or moved > 0, "the 3-rank case wants boxes that change owner (simple -> default partition)"
    # surviving boxes kept their data (also those that changed owner), new boxes are zero
    assert_state_equal(orc, mg, variables=(1, 2, 3, 4, 5), ghosts_of=())
    # ---- both go on: right-hand side of the new boxes, more cycles ----
    new_ids = [(lvl, id_) for lvl, id_ in P.all_ids(orc) if id_ > n_old]
    for m in (orc, mg):
        set_rhs(m, new_ids)
        # a standalone V-cycle refreshes the ghost cells of its finest level only, and the refinement-boundary interpolation
        # there reads ghost cells of the level below (src/m_ghost_cells.f90:500-577): after a change of the tree both sides
        # bring every level's ghost cells up to date, as an application does after copying its data into a new tree
        m.mg_fill_ghost_cells(P.mg_iphi)
    h_b = (P.run_cycles(orc, 3, False), P.run_cycles(mg, 3, False))
    assert h_b[0] == h_b[1], h_b
    assert_state_equal(orc, mg)
    assert h_b[0][-1] < 0.05 * h_b[0][0]


def test_update_tree_keeps_device_resident_data(oracle, product):
    """resident mode: nothing but the tree description crosses the boundary; the data is found again on the device"""
    mg = new_mg(product.Mg, gpu_resident=True)
    orc = new_mg(oracle.OracleMg)
    for m in (orc, mg):
        refine_level(m, 1)
        finish_tree(m, 2)
        m.mg_allocate_storage()
        set_rhs(m)
    for iv in (P.mg_iphi, P.mg_irhs, P.mg_num_vars + 1):
        mg.mg_gpu_push(iv)
    h = (P.run_cycles(orc, 2, False), P.run_cycles(mg, 2, False))
    assert h[0] == h[1]
    h2d_before = mg.stat("bytes_h2d")
    refine_level(mg, 2)
    finish_tree(mg, 3)
    mg.mg_gpu_update_tree()
    assert mg.stat("bytes_h2d") == h2d_before            # no variable went host -> device
    for iv in (P.mg_iphi, P.mg_irhs, P.mg_iold, P.mg_ires, P.mg_num_vars + 1):
        mg.mg_gpu_pull(iv, with_ghosts=False)
    for lvl, id_ in P.all_ids(orc):
        for iv in (1, 2, 3, 4, 5):
            a = orc.cc(id_)[iv - 1][1:-1, 1:-1, 1:-1]
            b = mg.cc(id_)[iv - 1][1:-1, 1:-1, 1:-1]
            assert np.array_equal(a, b), (lvl, id_, iv)
    for lvl, id_ in P.all_ids(mg):
        if id_ > orc.n_boxes:
            assert not mg.cc(id_)[:4, 1:-1, 1:-1, 1:-1].any()


@pytest.mark.parametrize("n_cpu", [2, 3])
def test_fresh_three_level_tree_on_ranks(oracle, product, n_cpu):
    """control for the test above: the same three-level tree built in one go (no regrid), random-free state"""
    kw = {"n_cpu": n_cpu}
    orc = new_mg(oracle.OracleMg, **kw)
    mg = new_mg(functools.partial(product.MgGroup, n_cpu))
    for m in (orc, mg):
        refine_level(m, 1)
        refine_level(m, 2)
        finish_tree(m, 3)
        m.mg_allocate_storage()
        set_rhs(m)
        for lvl, id_ in P.all_ids(m):
            X, Y, Z = P.cell_coords(m, id_, lvl)
            m.cc(id_)[P.mg_iphi - 1] = 0.3 * np.sin(1.7 * X) * np.cos(2.9 * Y + 0.2) + 0.1 * Z
        m.mg_fill_ghost_cells(P.mg_iphi)
    h = (P.run_cycles(orc, 3, False), P.run_cycles(mg, 3, False))
    assert h[0] == h[1], h
    assert_state_equal(orc, mg)
