"""CPU tests of the product's per-cell device code, compiled for the host (tests/device_host/harness.cpp).

The single-CTA tail of the cycle has a fast path for "simple" levels (power-of-two boxes whose faces are constant
physical boundaries or local neighbours, omg_ops.cuh simple_*).  The same headers the kernels are built from are
compiled with g++ -ffp-contract=off here, and the fast functions must reproduce the general ones -- the path the GPU
parity tests pin to the oracle -- bit for bit on random data: smoothing, residual, coarse right-hand side and
correction + prolongation, one or several boxes, every boundary-condition type."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "device_host", "harness.cpp")
LIB = os.path.join(HERE, "device_host", "libdevhost.so")
CSRC = os.path.join(os.path.dirname(HERE), "octree-mg_b200", "csrc")


@pytest.fixture(scope="module")
def devhost():
    deps = [SRC] + [os.path.join(CSRC, f) for f in ("omg_device.cuh", "omg_ops.cuh")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-w", "-I" + cuda_inc,
                               SRC, "-o", LIB])
    L = C.CDLL(LIB)
    L.dh_compare_simple.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    L.dh_compare_prolong.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_ulonglong]
    L.dh_ddiv.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    L.dh_recip6.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.dh_coeff_values.argtypes = [C.c_void_p] * 4 + [C.c_double] * 3 + [C.c_void_p]
    return L


BC_SETS = [[-10] * 6, [-11] * 6, [-12, -12, -10, -10, -11, -11], [-10, -11, -12, -10, -11, -12]]
BC_VALUES = [0.3, -0.7, 1.1, 0.0, 2.0, -1.5]
SHAPES = [(1, 1, 1), (2, 1, 1), (1, 2, 1), (1, 1, 2), (2, 2, 2), (3, 2, 1)]


def six(ctype, values):
    return (ctype * 6)(*values)


@pytest.mark.parametrize("planes", [0, 1])           # boundary values: constants / stored planes (callbacks)
@pytest.mark.parametrize("types", BC_SETS)
@pytest.mark.parametrize("nc", [2, 4, 8, 16])
def test_simple_level_functions_match_general_path(devhost, types, nc, planes):
    devhost.dh_set_planes(planes)
    for shape in SHAPES:
        for op, lam in ((1, 0.0), (2, 0.0), (3, 2.5), (4, 0.75), (5, 1.25)):   # mg_laplacian ... mg_ahelmholtz
            bad = devhost.dh_compare_simple(op, nc, *shape, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 7 * nc + op)
            assert bad == 0, (types, nc, shape, op)
    devhost.dh_set_planes(0)


@pytest.mark.parametrize("planes", [0, 1])
@pytest.mark.parametrize("types", BC_SETS)
@pytest.mark.parametrize("nc", [2, 4, 8, 16])
def test_simple_correction_and_prolongation_match_general_path(devhost, types, nc, planes):
    devhost.dh_set_planes(planes)
    for shape in SHAPES:                               # children = octants of parents of the same box size
        assert devhost.dh_compare_prolong(nc, *shape, 0, six(C.c_int, types), six(C.c_double, BC_VALUES), nc) == 0, shape
    if nc >= 4:                                        # single-box coarsening: the coarse box has half the size
        assert devhost.dh_compare_prolong(nc, 1, 1, 1, 1, six(C.c_int, types), six(C.c_double, BC_VALUES), nc) == 0
    devhost.dh_set_planes(0)


def test_batched_division_logic_on_host(devhost):
    """ddiv_batch with an exact-reciprocal seed in place of MUFU.RCP64H: the Newton / residual-correction sequence and
    the range test that sends a batch to the plain division (the device version is tested in test_gpu_steps.py)"""
    rng = np.random.default_rng(11)
    n = 6 * 20000
    a = rng.uniform(-4, 4, n) * 2.0 ** rng.integers(-1060, 1020, n)
    b = rng.uniform(1, 2, n) * 2.0 ** rng.integers(-1060, 1020, n) * rng.choice([-1.0, 1.0], n)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, 1.0, 3.0])
    k = len(special)
    a[:k * k] = np.repeat(special, k)
    b[:k * k] = np.tile(special, k)
    out = np.empty(n)
    devhost.dh_ddiv(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), n)
    with np.errstate(all="ignore"):
        ref = a / b
    same = (out.view(np.uint64) == ref.view(np.uint64)) | (np.isnan(out) & np.isnan(ref))
    assert same.all(), (a[~same][:5], b[~same][:5], out[~same][:5], ref[~same][:5])


def test_tolerance_mode_batch_inversion_is_accurate_to_a_few_ulp(devhost):
    """option coeff_math = 1 (NOT bitwise): six reciprocals from one division + 15 multiplications; each within a few ulp of
    1 / d, the per-cell operator and smoother values within 1e-14 relative of the exact ones; products outside the normal
    range fall back to exact divisions"""
    rng = np.random.default_rng(5)
    d = np.exp(rng.uniform(np.log(1e-6), np.log(1e6), 6 * 4000))
    out = np.empty_like(d)
    assert devhost.dh_recip6(d.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), d.size) == 0
    rel = np.abs(out * d - 1.0)
    assert rel.max() < 8 * np.finfo(float).eps, rel.max()
    big = np.full(6, 1e60)
    assert devhost.dh_recip6(big.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), 6) == 1 and np.array_equal(out[:6], 1.0 / big)
    idr2 = np.array([1024.0, 1024.0, 4096.0])
    for _ in range(200):
        u = rng.normal(size=6); a0 = np.full(6, np.exp(rng.uniform(-2, 2))); a = np.exp(rng.uniform(-2, 2, 6))
        res = np.zeros(4)
        devhost.dh_coeff_values(*(x.ctypes.data_as(C.c_void_p) for x in (idr2, u, a0, a)), float(rng.normal()), float(rng.normal()), 1.5,
                                res.ctypes.data_as(C.c_void_p))
        scale = 2 * idr2.sum() * a.max() * (np.abs(u).max() + 1)
        assert abs(res[0] - res[1]) <= 1e-14 * scale and abs(res[2] - res[3]) <= 1e-14 * (np.abs(u).max() + 1), res


BC_OF = {"dirichlet0": ([-10] * 6, [0.0] * 6), "neumann0": ([-11] * 6, [0.0] * 6),
         "mixed": ([-10] + [-11] * 5, [0.25] + [0.5] * 5), "continuous_x": ([-12, -12, -10, -10, -10, -10], [0.0, 0.0, 1.0, 1.0, 1.0, 1.0])}


@pytest.mark.parametrize("bc", sorted(BC_OF))
@pytest.mark.parametrize("op,lam", [(1, 0.0), (3, 1.75)])
def test_general_device_path_matches_oracle_on_one_level(devhost, oracle, op, lam, bc):
    """the device's general per-cell path (do_smooth, do_resid: colour-split layout, ghost values derived on the fly)
    against the oracle's smooth_boxes / max_residual_lvl (ghost cells stored and refilled, src/m_multigrid.f90:404-436)
    on the 2 x 2 x 2 boxes of one level with random data: same bits, on the CPU"""
    import problems as P
    devhost.dh_level_op.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double),
                                                   C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    nc = 8
    orc, _ = P.uniform_problem(oracle.OracleMg, 16, nc, operator_type=op, bc=bc, lam=lam)
    lvl = orc.highest_lvl
    ids = orc.ids(lvl)
    assert len(ids) == 8
    rng = np.random.default_rng(5)
    pos = {}
    for id_ in ids:
        ix = orc.box_info(id_)["ix"]
        pos[id_] = ((ix[2] - 1) * 2 + (ix[1] - 1)) * 2 + (ix[0] - 1)
        cc = orc.cc(id_, nc)
        cc[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc))
        cc[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc)) * 100.0
    orc.mg_fill_ghost_cells_lvl(lvl, P.mg_iphi)

    def gather(iv):
        out = np.empty((8, nc, nc, nc))
        for id_ in ids:
            out[pos[id_]] = orc.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
        return out

    types, values = BC_OF[bc]
    dr = (C.c_double * 3)(*orc.dr(lvl))
    phi, rhs = gather(P.mg_iphi), gather(P.mg_irhs)
    res = np.zeros_like(phi)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)          # noqa: E731
    for sweep in range(2):
        devhost.dh_level_op(0, op, nc, 2, 2, 2, six(C.c_int, types), six(C.c_double, values), lam, dr, vp(phi), vp(rhs), vp(res), None)
        orc.smooth_boxes(lvl, 1)
        assert np.array_equal(phi, gather(P.mg_iphi)), "GSRB sweep %d differs" % sweep
    m = C.c_double(0.0)
    devhost.dh_level_op(1, op, nc, 2, 2, 2, six(C.c_int, types), six(C.c_double, values), lam, dr, vp(phi), vp(rhs), vp(res), C.byref(m))
    assert orc.max_residual_lvl(lvl) == m.value
    assert np.array_equal(res, gather(P.mg_ires))


@pytest.mark.parametrize("bc", sorted(BC_OF))
@pytest.mark.parametrize("op,lam", [(1, 0.0), (3, 1.75)])
def test_device_lexicographic_smoother_matches_oracle_on_one_level(devhost, oracle, op, lam, bc):
    """mg_smoother_gs, the reference's default smoother: the device sweeps hyperplanes i + j + k = w against a snapshot of
    the ghost faces (do_gs_snapshot + do_gs_sweep); the oracle runs the reference's lexicographic k, j, i loop per box after
    a ghost fill (box_gs_lpl with redblack = .false., src/m_laplacian.f90:99-111) -- same bits after every sweep"""
    import problems as P
    devhost.dh_level_op.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double),
                                                   C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    nc = 8
    orc, _ = P.uniform_problem(oracle.OracleMg, 16, nc, operator_type=op, bc=bc, lam=lam, smoother=P.mg_smoother_gs)
    lvl = orc.highest_lvl
    ids = orc.ids(lvl)
    rng = np.random.default_rng(17)
    pos = {}
    for id_ in ids:
        ix = orc.box_info(id_)["ix"]
        pos[id_] = ((ix[2] - 1) * 2 + (ix[1] - 1)) * 2 + (ix[0] - 1)
        cc = orc.cc(id_, nc)
        cc[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc))
        cc[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc)) * 100.0
    orc.mg_fill_ghost_cells_lvl(lvl, P.mg_iphi)

    def gather(iv):
        out = np.empty((8, nc, nc, nc))
        for id_ in ids:
            out[pos[id_]] = orc.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
        return out

    types, values = BC_OF[bc]
    dr = (C.c_double * 3)(*orc.dr(lvl))
    phi, rhs = gather(P.mg_iphi), gather(P.mg_irhs)
    res = np.zeros_like(phi)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)          # noqa: E731
    for sweep in range(3):
        devhost.dh_level_op(2, op, nc, 2, 2, 2, six(C.c_int, types), six(C.c_double, values), lam, dr, vp(phi), vp(rhs), vp(res), None)
        orc.smooth_boxes(lvl, 1)
        assert np.array_equal(phi, gather(P.mg_iphi)), "lexicographic sweep %d differs" % sweep


@pytest.mark.parametrize("bc", ["dirichlet0", "mixed"])
@pytest.mark.parametrize("op,lam", [(2, 0.0), (4, 1.75), (5, 0.5)])
def test_variable_coefficient_device_path_matches_oracle_on_one_level(devhost, oracle, op, lam, bc):
    """mg_vlaplacian / mg_vhelmholtz / mg_ahelmholtz: the device's general path (harmonic-mean face coefficients through the
    batched division, eps ghost cells as stored faces) against the oracle's box_gs_v* / box_v* (src/m_vlaplacian.f90:51-189,
    m_vhelmholtz.f90:61-205, m_ahelmholtz.f90:69-237) on 2 x 2 x 2 boxes with random phi and rhs: same bits, on the CPU"""
    import problems as P
    devhost.dh_level_op_vc.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double),
                                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_double)]
    nc = 8
    orc, _ = P.uniform_problem(oracle.OracleMg, 16, nc, operator_type=op, bc=bc, lam=lam)
    ne = P.n_eps(op)
    lvl = orc.highest_lvl
    ids = orc.ids(lvl)
    rng = np.random.default_rng(23)
    pos = {}
    eps_cc = np.empty((ne, 8, nc + 2, nc + 2, nc + 2))
    for id_ in ids:
        ix = orc.box_info(id_)["ix"]
        pos[id_] = ((ix[2] - 1) * 2 + (ix[1] - 1)) * 2 + (ix[0] - 1)
        cc = orc.cc(id_, nc)
        cc[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc))
        cc[P.mg_irhs - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc)) * 100.0
        for d in range(ne):
            eps_cc[d, pos[id_]] = cc[P.mg_iveps - 1 + d]
    orc.mg_fill_ghost_cells_lvl(lvl, P.mg_iphi)

    def gather(iv):
        out = np.empty((8, nc, nc, nc))
        for id_ in ids:
            out[pos[id_]] = orc.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
        return out

    types, values = BC_OF[bc]
    dr = (C.c_double * 3)(*orc.dr(lvl))
    phi, rhs = gather(P.mg_iphi), gather(P.mg_irhs)
    res = np.zeros_like(phi)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)          # noqa: E731
    args = (op, nc, 2, 2, 2, six(C.c_int, types), six(C.c_double, values), lam, dr, vp(phi), vp(rhs), vp(res), vp(eps_cc), ne)
    for sweep in range(2):
        devhost.dh_level_op_vc(0, *args, None)
        orc.smooth_boxes(lvl, 1)
        assert np.array_equal(phi, gather(P.mg_iphi)), "GSRB sweep %d differs" % sweep
    m = C.c_double(0.0)
    devhost.dh_level_op_vc(1, *args, C.byref(m))
    assert orc.max_residual_lvl(lvl) == m.value
    assert np.array_equal(res, gather(P.mg_ires))


@pytest.mark.parametrize("simple", [0, 1])
@pytest.mark.parametrize("bc", sorted(BC_OF))
@pytest.mark.parametrize("op,lam", [(1, 0.0), (3, 1.75)])
def test_device_transfer_operators_match_oracle(devhost, oracle, op, lam, bc, simple):
    """update_coarse (residual, restriction, coarse right-hand side, old = phi) and correct_children (correction +
    prolongation) of the device code -- general and fast variants -- against the oracle on two levels (one parent box,
    its eight children) with random data: same bits, on the CPU"""
    import problems as P
    devhost.dh_two_level.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double)] + [C.c_void_p] * 8
    nc = 8
    orc, _ = P.uniform_problem(oracle.OracleMg, 16, nc, operator_type=op, bc=bc, lam=lam)
    fine, coarse = orc.highest_lvl, orc.highest_lvl - 1
    fids, (cid,) = orc.ids(fine), orc.ids(coarse)
    assert len(fids) == 8 and orc.box_size_lvl(coarse) == nc
    rng = np.random.default_rng(9)
    pos = {}
    for id_ in fids:
        ix = orc.box_info(id_)["ix"]
        pos[id_] = ((ix[2] - 1) * 2 + (ix[1] - 1)) * 2 + (ix[0] - 1)
    for id_ in list(fids) + [cid]:
        cc = orc.cc(id_, nc)
        for iv in (P.mg_iphi, P.mg_irhs, P.mg_iold, P.mg_ires):
            cc[iv - 1, 1:-1, 1:-1, 1:-1] = rng.standard_normal((nc, nc, nc))
    orc.mg_fill_ghost_cells_lvl(fine, P.mg_iphi)
    orc.mg_fill_ghost_cells_lvl(coarse, P.mg_iphi)

    def gather(ids, iv):
        out = np.empty((len(ids), nc, nc, nc))
        for id_ in ids:
            out[pos.get(id_, 0)] = orc.cc(id_, nc)[iv - 1, 1:-1, 1:-1, 1:-1]
        return out

    ivs = (P.mg_iphi, P.mg_irhs, P.mg_iold, P.mg_ires)
    f = [gather(fids, iv) for iv in ivs]
    c = [gather([cid], iv) for iv in ivs]
    types, values = BC_OF[bc]
    dr = (C.c_double * 3)(*orc.dr(fine))
    vp = lambda a: a.ctypes.data_as(C.c_void_p)          # noqa: E731
    call = lambda what: devhost.dh_two_level(what, simple, op, nc, six(C.c_int, types), six(C.c_double, values), lam, dr,   # noqa: E731
                                             *[vp(a) for a in f], *[vp(a) for a in c])
    call(0)
    orc.update_coarse(fine)
    assert np.array_equal(f[3], gather(fids, P.mg_ires)), "fine residual"
    for k, iv in enumerate(ivs):
        assert np.array_equal(c[k], gather([cid], iv)), "coarse variable %d after update_coarse" % iv
    # a coarse "solve": new phi_c, ghost cells refilled as smooth_boxes would leave them
    new_phi = rng.standard_normal((nc, nc, nc))
    orc.cc(cid, nc)[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1] = new_phi
    orc.mg_fill_ghost_cells_lvl(coarse, P.mg_iphi)
    c[0][0] = new_phi
    call(1)
    orc.correct_children(coarse)
    assert np.array_equal(f[0], gather(fids, P.mg_iphi)), "fine phi after correct_children"
    assert np.array_equal(c[3], gather([cid], P.mg_ires)), "coarse correction"


@pytest.mark.parametrize("types", BC_SETS[:3])
@pytest.mark.parametrize("nc,shape,tx,ty,S", [(8, (2, 2, 2), 8, 8, 4), (8, (3, 2, 2), 16, 8, 4), (4, (4, 4, 2), 8, 8, 2),
                                              (8, (2, 2, 1), 4, 4, 4), (8, (2, 2, 2), 16, 16, 6), (16, (2, 1, 1), 16, 8, 4),
                                              (8, (2, 2, 2), 6, 10, 3)])
def test_temporal_blocking_prototype_matches_separate_half_sweeps(devhost, types, nc, shape, tx, ty, S):
    """next-round groundwork (harness.cpp dh_fused_proto): S GSRB half-sweeps applied in one
    z-streaming pass over xy tiles with an S-cell halo, in place, give the bits of S separate half-sweeps of the
    general device path -- for every tiling, S = 2..6 and every boundary-condition type"""
    devhost.dh_fused_proto.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_ulonglong]
    assert devhost.dh_fused_proto(nc, *shape, tx, ty, S, six(C.c_int, types), six(C.c_double, BC_VALUES), 5) == 0
