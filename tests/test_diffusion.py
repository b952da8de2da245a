"""Implicit diffusion solvers (src/m_diffusion.f90:19-173).

CPU part: the oracle's restatement is pinned by a closed form -- for a discrete eigenfunction of the
cell-centred Laplacian with homogeneous Dirichlet faces one backward-Euler (order 1) or Crank-Nicolson
(order 2) step multiplies the field by a known factor.  GPU part: the product (host mirror + C ABI,
`omg_diffusion_set_rhs` between `omg_apply_op`, `omg_fas_fmg` and `omg_fas_vcycle`) against the oracle, bitwise."""
import functools

import numpy as np
import pytest

import problems as P
from compare import assert_state_equal

MAX_RES = 1e-9


def amplification(n3, dt, coeff, order):
    lam = P.diffusion_eigenvalue(n3) * dt * coeff
    return 1.0 / (1.0 + lam) if order == 1 else (1.0 - 0.5 * lam) / (1.0 + 0.5 * lam)


def finest_ratio_error(mg, factor):
    err = 0.0
    lvl = mg.highest_lvl
    for id_ in mg.ids(lvl):
        X, Y, Z = P.cell_coords(mg, id_, lvl)
        want = factor * P.diffusion_mode(X, Y, Z)[1:-1, 1:-1, 1:-1]
        err = max(err, float(np.abs(mg.cc(id_)[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1] - want).max()))
    return err


@pytest.mark.parametrize("order", [1, 2])
def test_oracle_constant_coefficient_step_is_exact(oracle, order):
    n, dt, coeff = 32, 2.0e-3, 1.5
    orc = P.diffusion_problem(oracle.OracleMg, n, 8)
    res, n_vc = orc.diffusion_solve(dt, coeff, order, MAX_RES)
    assert res <= MAX_RES and 0 <= n_vc <= 10
    assert finest_ratio_error(orc, amplification([n] * 3, dt, coeff, order)) < 1e-10


def test_oracle_set_rhs_touches_leaves_of_positive_levels_only(oracle):
    orc = P.diffusion_problem(oracle.OracleMg, 16, 8)
    before = {id_: orc.cc(id_).copy() for _, id_ in P.all_ids(orc)}
    orc.diffusion_set_rhs(-3.0, 0.5)
    for lvl, id_ in P.all_ids(orc):
        cc, old = orc.cc(id_), before[id_]
        leaf = lvl >= 1 and id_ in set(orc.leaves(lvl))
        want = old[P.mg_irhs - 1].copy()
        if leaf:
            want[1:-1, 1:-1, 1:-1] = (-3.0 * old[P.mg_iphi - 1] + 0.5 * old[P.mg_irhs - 1])[1:-1, 1:-1, 1:-1]
        assert np.array_equal(cc[P.mg_irhs - 1], want)
        assert np.array_equal(cc[P.mg_iphi - 1], old[P.mg_iphi - 1])


def test_oracle_rejects_bad_order(oracle):
    orc = P.diffusion_problem(oracle.OracleMg, 16, 8)
    with pytest.raises(Exception, match="order should be 1 or 2"):
        orc.diffusion_solve(1e-3, 1.0, 3, MAX_RES)


def test_oracle_variable_coefficient_reduces_to_constant(oracle):
    """eps = const: diffusion_solve_vcoeff(dt) solves the same system as diffusion_solve(dt, eps)"""
    n, dt, coeff = 16, 1.0e-3, 2.0
    a = P.diffusion_problem(oracle.OracleMg, n, 8)
    b = P.diffusion_problem(oracle.OracleMg, n, 8, operator_type=P.mg_vhelmholtz)
    for _, id_ in P.all_ids(b):
        b.cc(id_)[P.mg_iveps - 1] = coeff
    a.diffusion_solve(dt, coeff, 1, 1e-11)
    b.diffusion_solve_vcoeff(dt, 1, 1e-11)
    f = amplification([n] * 3, dt, coeff, 1)
    assert finest_ratio_error(a, f) < 1e-10 and finest_ratio_error(b, f) < 1e-10


# ---------------------------------------------------------------------------------------------------------------
# GPU parity
# ---------------------------------------------------------------------------------------------------------------
def solve(mg, op, dt, order, max_res=MAX_RES):
    if op == P.mg_helmholtz:
        return mg.diffusion_solve(dt, 1.5, order, max_res)
    if op == P.mg_vhelmholtz:
        return mg.diffusion_solve_vcoeff(dt, order, max_res)
    return mg.diffusion_solve_acoeff(dt, order, max_res)


@pytest.mark.gpu
@pytest.mark.parametrize("order", [1, 2])
@pytest.mark.parametrize("op", [P.mg_helmholtz, P.mg_vhelmholtz, P.mg_ahelmholtz])
@pytest.mark.parametrize("n,box", [(32, 8), (32, 16)])
def test_gpu_diffusion_uniform(oracle, product, op, order, n, box):
    orc = P.diffusion_problem(oracle.OracleMg, n, box, operator_type=op)
    mg = P.diffusion_problem(product.Mg, n, box, operator_type=op)
    for step in range(2):                      # two time steps: the second starts from the device's own result
        assert solve(orc, op, 2.0e-3, order) == solve(mg, op, 2.0e-3, order)
    assert_state_equal(orc, mg)
    if op == P.mg_helmholtz:
        assert finest_ratio_error(mg, amplification([n] * 3, 2.0e-3, 1.5, order) ** 2) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("order", [1, 2])
def test_gpu_diffusion_refined(oracle, product, order):
    orc = P.diffusion_problem(oracle.OracleMg, 16, 8, amr_levels=3)
    mg = P.diffusion_problem(product.Mg, 16, 8, amr_levels=3)
    # |rhs| ~ 1/dt = 1e3 and the refinement boundaries slow the cycle down: ask for 9 orders, not 12, within max_its
    assert solve(orc, P.mg_helmholtz, 1.0e-3, order, 1e-6) == solve(mg, P.mg_helmholtz, 1.0e-3, order, 1e-6)
    assert_state_equal(orc, mg)


@pytest.mark.gpu
def test_gpu_diffusion_no_convergence_is_an_error(oracle, product):
    """max_its = 10 V-cycles after the FMG cycle (m_diffusion.f90:25,49-58): both sides stop with the same error"""
    orc = P.diffusion_problem(oracle.OracleMg, 16, 8, amr_levels=3)
    mg = P.diffusion_problem(product.Mg, 16, 8, amr_levels=3)
    with pytest.raises(Exception, match="no convergence"):
        orc.diffusion_solve(1.0e-3, 1.5, 1, 1e-30)
    with pytest.raises(product.MgError, match="no convergence"):
        mg.diffusion_solve(1.0e-3, 1.5, 1, 1e-30)
    assert_state_equal(orc, mg)


@pytest.mark.gpu
def test_gpu_diffusion_resident_and_switching_operator(oracle, product):
    """device-resident: push once, several steps, pull once; the mg_t was set up for a Laplacian (the operator is
    switched by diffusion_solve itself, :29-30)"""
    orc = P.diffusion_problem(oracle.OracleMg, 32, 16, operator_type=P.mg_laplacian)
    mg = P.diffusion_problem(product.Mg, 32, 16, operator_type=P.mg_laplacian, gpu_resident=True)
    mg.mg_gpu_push(P.mg_iphi)
    mg.mg_gpu_push(P.mg_irhs)
    h2d = mg.stat("bytes_h2d")
    for order in (1, 2, 1):
        assert orc.diffusion_solve(1.0e-3, 0.7, order, MAX_RES) == mg.diffusion_solve(1.0e-3, 0.7, order, MAX_RES)
    assert mg.stat("bytes_h2d") - h2d < 1 << 20          # parameters and tables only: no field crossed PCIe
    for iv in (P.mg_iphi, P.mg_irhs, P.mg_iold, P.mg_ires):
        mg.mg_gpu_pull(iv, iv in (P.mg_iphi, P.mg_iold))
    assert_state_equal(orc, mg)


@pytest.mark.gpu
@pytest.mark.parametrize("n_cpu", [2, 4])
def test_gpu_diffusion_ranks(oracle, product, n_cpu):
    orc = P.diffusion_problem(oracle.OracleMg, (64, 32, 32), 16, n_cpu=n_cpu)
    mg = P.diffusion_problem(functools.partial(product.MgGroup, n_cpu), (64, 32, 32), 16)
    for order in (1, 2):
        assert orc.diffusion_solve(1.0e-3, 1.0, order, MAX_RES) == mg.diffusion_solve(1.0e-3, 1.0, order, MAX_RES)
    assert_state_equal(orc, mg)


@pytest.mark.gpu
def test_gpu_diffusion_errors(product):
    mg = P.diffusion_problem(product.Mg, 16, 8)
    with pytest.raises(product.MgError, match="order should be 1 or 2"):
        mg.diffusion_solve(1e-3, 1.0, 0, MAX_RES)
    with pytest.raises(product.MgError, match="n_extra_vars"):
        mg.diffusion_solve_vcoeff(1e-3, 1, MAX_RES)       # allocated without mg_iveps (m_vhelmholtz.f90:23-27)
