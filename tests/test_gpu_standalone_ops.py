"""Stand-alone level operations on arbitrary variables, single- and multi-rank:
mg_fill_ghost_cells_lvl for variables >= 5 (src/m_ghost_cells.f90:131-175: same-level neighbours, physical boundaries
with the variable's own constants, refinement boundaries, faces owned by other ranks), mg_prolong of any variable
(src/m_prolong.f90:51-85, parents on other ranks included) and mg_restrict_lvl (src/m_restrict.f90:83-114).
The free-space front end (m_free_space.f90:185-199) drives exactly this sequence.  Everything bitwise against the
oracle, ghost faces included."""
import functools

import numpy as np
import pytest

import problems as P
from compare import assert_state_equal

pytestmark = pytest.mark.gpu

IV = P.mg_num_vars + 1          # the extra variable
IW = P.mg_num_vars + 2          # a second one (prolongation target)


def field(X, Y, Z):
    return np.sin(3.1 * X + 0.2) * np.cos(2.3 * Y) + 0.5 * Z * Z - 0.3 * X * Y + 0.7


def build(cls, amr_levels, n=32, box=8, **kw):
    mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, n_extra_vars=2, **kw)
    kinds = [P.mg_bc_dirichlet, P.mg_bc_neumann, P.mg_bc_continuous]
    for nb in range(1, 7):
        mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.125 * nb)
        mg.set_bc(nb, IV, kinds[nb % 3], 0.3 * nb - 0.4)
        mg.set_bc(nb, IW, kinds[(nb + 1) % 3], 0.1 * nb)
    mg.mg_set_methods()
    if amr_levels:
        P.build_amr_tree(mg, amr_levels, [n] * 3, box, [1.0 / n] * 3)
    else:
        mg.mg_build_rectangle([n] * 3, box, [1.0 / n] * 3, [0.0] * 3, [False] * 3)
        mg.mg_load_balance()
        mg.mg_allocate_storage()
    for lvl, id_ in P.all_ids(mg):
        X, Y, Z = P.cell_coords(mg, id_, lvl)
        cc = mg.cc(id_)
        cc[IV - 1] = field(X, Y, Z) * (1.0 + 0.1 * lvl)
        cc[IW - 1] = 0.01 * field(Z, X, Y)
        cc[P.mg_iphi - 1] = field(Y, Z, X)
    return mg


def pair(oracle, product, n_cpu, amr_levels, **kw):
    if n_cpu == 1:
        return build(oracle.OracleMg, amr_levels, **kw), build(product.Mg, amr_levels, **kw)
    return build(oracle.OracleMg, amr_levels, n_cpu=n_cpu, **kw), build(functools.partial(product.MgGroup, n_cpu), amr_levels, **kw)


@pytest.mark.parametrize("amr_levels", [0, 3])
@pytest.mark.parametrize("n_cpu", [1, 2, 3, 5])
def test_fill_ghost_cells_of_an_extra_variable(oracle, product, n_cpu, amr_levels):
    orc, mg = pair(oracle, product, n_cpu, amr_levels)
    for m in (orc, mg):
        for lvl in range(m.lowest_lvl, m.highest_lvl + 1):      # coarse to fine: refinement boundaries read coarse ghosts
            m.mg_fill_ghost_cells_lvl(lvl, IV)
    assert_state_equal(orc, mg, variables=(IV,), ghosts_of=(IV,))


@pytest.mark.parametrize("amr_levels", [0, 3])
@pytest.mark.parametrize("n_cpu", [1, 3, 4])
def test_restrict_prolong_fill_sequence_of_the_free_space_front_end(oracle, product, n_cpu, amr_levels):
    """restrict a variable down, then prolong it back up level by level with a ghost fill after each prolongation
    (src/m_free_space.f90:185-199), for an extra variable and for phi"""
    orc, mg = pair(oracle, product, n_cpu, amr_levels)
    for iv in (IV, P.mg_iphi):
        for m in (orc, mg):
            for lvl in range(m.highest_lvl, m.lowest_lvl, -1):
                m.mg_restrict_lvl(iv, lvl)
            m.mg_fill_ghost_cells_lvl(m.lowest_lvl, iv)
            for lvl in range(m.lowest_lvl, m.highest_lvl):
                m.mg_prolong(lvl, iv, iv, False)
                m.mg_fill_ghost_cells_lvl(lvl + 1, iv)
        assert_state_equal(orc, mg, variables=(iv,), ghosts_of=(iv,))


@pytest.mark.parametrize("n_cpu", [1, 3])
def test_prolong_into_another_variable_with_add(oracle, product, n_cpu):
    orc, mg = pair(oracle, product, n_cpu, 3)
    for m in (orc, mg):
        for lvl in range(m.lowest_lvl, m.highest_lvl + 1):
            m.mg_fill_ghost_cells_lvl(lvl, IV)
            m.mg_fill_ghost_cells_lvl(lvl, P.mg_iphi)     # the device derives phi's ghost cells: make the oracle's stored ones consistent
        for lvl in range(m.lowest_lvl, m.highest_lvl):
            m.mg_prolong(lvl, IV, IW, True)
            m.mg_prolong(lvl, P.mg_iphi, P.mg_irhs, False)
    assert_state_equal(orc, mg, variables=(IW, P.mg_irhs, IV), ghosts_of=(IV,))
