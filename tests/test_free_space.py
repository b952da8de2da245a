"""Free-space front end: mg_poisson_free_3d, src/m_free_space.f90:36-214, on the problem of
the reference's tests/test_free_space.f90 (a Gaussian charge, solution erf(r / sigma) / (4 pi r)).

CPU: the oracle's restatement (oracle/free_space.inc: the order-8 interpolating-scaling-function kernel of 1/r of the
vendored BigDFT solver + a direct separable convolution) is pinned by the analytic solution -- with the FFT level equal
to the finest level the free-space solve alone must reproduce erf(r/sigma)/(4 pi r) to the accuracy of the 89-term
Gaussian quadrature (1e-8), which no wrong kernel, scaling or index convention survives.
GPU: the product (host kernel construction + cuFFT convolution + device-side boundary planes) against the oracle.  The two
convolutions add the same terms in a different order, so the comparison uses the north-star tolerances (1e-10 relative on
the residuals, 1e-12 * |phi| on the solution), not bit equality."""
import functools

import numpy as np
import pytest

import problems as P

SIGMA = 0.1
R0 = (0.5, 0.5, 0.5)


def erf_vec(x):
    import math
    return np.vectorize(math.erf)(x)


def solution(X, Y, Z):      # tests/test_free_space.f90:99-111
    r = np.sqrt((X - R0[0]) ** 2 + (Y - R0[1]) ** 2 + (Z - R0[2]) ** 2)
    fac = 1 / (4 * np.pi)
    return np.where(r < np.sqrt(np.finfo(float).eps), 2 * fac / (np.sqrt(np.pi) * SIGMA), fac * erf_vec(r / SIGMA) / np.maximum(r, 1e-300))


def rhs(X, Y, Z):           # :113-121
    r2 = ((X - R0[0]) ** 2 + (Y - R0[1]) ** 2 + (Z - R0[2]) ** 2) / SIGMA ** 2
    return -1.0 / (SIGMA ** 3 * np.pi * np.sqrt(np.pi)) * np.exp(-r2)


def build(cls, n, box, amr_levels=0, **kw):
    mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, n_extra_vars=1, **kw)
    mg.mg_set_methods()
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    if amr_levels:
        P.build_amr_tree(mg, amr_levels, n3, box, [1.0 / n3[0]] * 3)
    else:
        mg.mg_build_rectangle(n3, box, [1.0 / n3[0]] * 3, [0.0] * 3, [False] * 3)
        mg.mg_load_balance()
        mg.mg_allocate_storage()
    for lvl, id_ in P.all_ids(mg):
        X, Y, Z = P.cell_coords(mg, id_, lvl)
        mg.cc(id_)[P.mg_irhs - 1] = rhs(X, Y, Z)
    return mg


def max_error(mg):
    err = 0.0
    for id_ in mg.ids(mg.highest_lvl):
        X, Y, Z = P.cell_coords(mg, id_, mg.highest_lvl)
        err = max(err, float(np.abs(mg.cc(id_)[P.mg_iphi - 1][1:-1, 1:-1, 1:-1] - solution(X, Y, Z)[1:-1, 1:-1, 1:-1]).max()))
    return err


def test_oracle_free_space_solve_reproduces_the_analytic_potential(oracle):
    """FFT level = finest level: no multigrid at all, the potential is the convolution itself.  The error falls with the
    order of the scaling functions (1.4e-5 at 32^3 -> 8e-8 at 64^3: a factor 170 ~ 2^7.4) down to the accuracy of the
    Gaussian quadrature (1e-8)"""
    errs = {}
    for n in (32, 64):
        mg = build(oracle.OracleMg, n, 16)
        mg.mg_poisson_free_3d(True, 1.0)
        assert mg.free_space_fft_lvl == mg.highest_lvl
        errs[n] = max_error(mg)
    assert errs[64] < 5e-7 and errs[32] > 50 * errs[64], errs


def test_oracle_free_space_multigrid_converges_to_the_discretisation_error(oracle):
    """tests/test_free_space.f90: FMG cycles with boundary values from the coarse FFT solve; the error settles at
    O(h^2) and the residual drops by more than an order of magnitude per cycle"""
    errs = {}
    for n in (32, 64):
        mg = build(oracle.OracleMg, n, 8)
        res = [mg.mg_poisson_free_3d(it == 0, 0.15) for it in range(3)]
        assert mg.free_space_fft_lvl < mg.highest_lvl
        assert res[1] < 0.1 * res[0] and res[2] < 0.1 * res[1], res
        errs[n] = max_error(mg)
    assert errs[64] < 0.35 * errs[32] and errs[64] < 3e-3, errs


def test_oracle_free_space_first_call_needs_a_rhs(oracle):
    mg = build(oracle.OracleMg, 16, 8)
    with pytest.raises(oracle.OracleError, match="first call requires new_rhs"):
        mg.mg_poisson_free_3d(False, 0.15)


CASES = [
    # n, box, amr levels, max_fft_frac, ranks
    (32, 8, 0, 0.15, 1),
    (32, 16, 0, 1.0, 1),            # FFT level = finest level
    (64, 16, 0, 0.15, 1),           # TMA-staged kernels above the FFT level
    ((32, 32, 64), 8, 0, 0.2, 1),   # non-cubic domain
    (32, 8, 3, 0.3, 1),             # AMR: FFT on the highest fully refined level or below
    (32, 8, 0, 0.15, 2),
    (32, 8, 0, 0.15, 4),
    (32, 8, 3, 0.3, 3),
]


@pytest.mark.gpu
@pytest.mark.parametrize("n,box,amr,frac,n_cpu", CASES)
def test_gpu_free_space_matches_the_oracle(oracle, product, n, box, amr, frac, n_cpu):
    orc = build(oracle.OracleMg, n, box, amr, **({"n_cpu": n_cpu} if n_cpu > 1 else {}))
    mg = build(product.Mg if n_cpu == 1 else functools.partial(product.MgGroup, n_cpu), n, box, amr)
    h_o = [orc.mg_poisson_free_3d(it == 0, frac) for it in range(3)]
    h_g = [mg.mg_poisson_free_3d(it == 0, frac) for it in range(3)]
    assert orc.free_space_fft_lvl == mg.free_space_fft_lvl
    # the rounding floor of a residual: phi carries ~1e-15 relative differences (two summation orders of the same
    # convolution), the operator multiplies them by 2 * sum(1 / dr^2)
    dr = mg.dr(mg.highest_lvl)
    floor = 64 * np.finfo(float).eps * 2.0 * float(np.sum(1.0 / np.asarray(dr) ** 2))
    for a, b in zip(h_o, h_g):
        assert abs(a - b) <= 1e-10 * abs(a) + floor, (h_o, h_g, floor)
    scale = 0.0
    diff = 0.0
    for lvl, id_ in P.all_ids(orc):
        a = orc.cc(id_)[P.mg_iphi - 1][1:-1, 1:-1, 1:-1]
        b = mg.cc(id_)[P.mg_iphi - 1][1:-1, 1:-1, 1:-1]
        scale = max(scale, float(np.abs(a).max()))
        diff = max(diff, float(np.abs(a - b).max()))
    assert diff <= 1e-12 * scale, (diff, scale)
    # V-cycles instead of FMG, and a second right-hand side on the same FFT grid
    for m in (orc, mg):
        for lvl, id_ in P.all_ids(m):
            m.cc(id_)[P.mg_irhs - 1] *= 0.5
    r_o = [orc.mg_poisson_free_3d(True, frac, fmgcycle=False), orc.mg_poisson_free_3d(False, frac, fmgcycle=False)]
    r_g = [mg.mg_poisson_free_3d(True, frac, fmgcycle=False), mg.mg_poisson_free_3d(False, frac, fmgcycle=False)]
    for a, b in zip(r_o, r_g):
        assert abs(a - b) <= 1e-10 * abs(a) + floor, (r_o, r_g, floor)


@pytest.mark.gpu
def test_gpu_free_space_errors(product):
    mg = build(product.Mg, 16, 8)
    with pytest.raises(product.MgError, match="first call requires new_rhs"):
        mg.mg_poisson_free_3d(False, 0.15)
    mg2 = product.Mg(P.mg_helmholtz, P.mg_smoother_gsrb)
    mg2.mg_set_methods()
    mg2.mg_build_rectangle([16] * 3, 8, [1.0 / 16] * 3, [0.0] * 3, [False] * 3)
    mg2.mg_load_balance()
    mg2.mg_allocate_storage()
    with pytest.raises(product.MgError, match="laplacian operator required"):
        mg2.mg_poisson_free_3d(True, 0.15)
