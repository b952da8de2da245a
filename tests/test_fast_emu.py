"""CPU tests of the TMA-staged kernels (octree-mg_b200/csrc/omg_fast.cu) through the CTA emulator
(tests/device_host/fast_emu.cpp).

g++ compiles the kernel source the GPU runs; 256 host threads stand in for a CTA, a TMA bulk copy is a memcpy whose
alignment, size, destination and mbarrier byte count are checked.  On random levels the fast kernels must reproduce the
general per-cell path (omg_ops.cuh do_*) bit for bit -- the path test_device_host.py pins to the oracle on the CPU and
the GPU parity tests pin on the device: GSRB half-sweeps (whole level and box lists), residual + maximum, coarse
right-hand side, residual fused with restriction, correction fused with prolongation; box size 16 (whole-box tiles)
and 32 (four slabs per box); Laplacian, Helmholtz and their variable-coefficient forms; every boundary-condition type
as a constant, as stored planes, and ghost planes (refinement / partition boundaries)."""
import ctypes as C
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
DH = os.path.join(HERE, "device_host")
LIB = os.path.join(DH, "libfastemu.so")
CSRC = os.path.join(os.path.dirname(HERE), "octree-mg_b200", "csrc")


@pytest.fixture(scope="module")
def emu():
    deps = [os.path.join(DH, "fast_emu.cpp"), os.path.join(DH, "harness.cpp")] + \
           [os.path.join(CSRC, f) for f in ("omg_device.cuh", "omg_ops.cuh", "omg_fast.cu", "omg_fast.h", "omg_tail.cuh", "omg_kernels.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
        # -Bsymbolic: the emulator's omgf::* must not be interposed by libomg_b200.so when other tests loaded it first
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-w", "-pthread", "-Wl,-Bsymbolic",
                               "-I" + cuda_inc, os.path.join(DH, "fast_emu.cpp"), "-o", LIB])
    L = C.CDLL(LIB)
    for f in (L.fe_one_level, L.fe_two_levels):
        f.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    L.dh_set_planes.argtypes = [C.c_int]
    return L


BC_SETS = [[-10] * 6, [-11] * 6, [-12, -12, -10, -10, -11, -11], [-10, -11, -12, -10, -11, -12]]
BC_VALUES = [0.3, -0.7, 1.1, 0.0, 2.0, -1.5]
OPS = [(1, 0.0), (2, 0.0), (3, 2.5), (4, 0.75), (5, 1.25)]      # mg_laplacian, mg_vlaplacian, mg_helmholtz, mg_vhelmholtz, mg_ahelmholtz


def six(ctype, values):
    return (ctype * 6)(*values)


@pytest.mark.parametrize("planes", [0, 1, 2])        # boundary constants / stored boundary planes / ghost planes
@pytest.mark.parametrize("op,lam", OPS)
@pytest.mark.parametrize("nc", [16, 32])
def test_fast_kernels_match_general_path_on_one_level(emu, nc, op, lam, planes):
    emu.dh_set_planes(planes)
    try:
        cases = list(zip(BC_SETS, [(2, 2, 1), (1, 2, 2), (2, 1, 2), (3, 1, 1)]))
        if nc == 32 or planes == 2:                  # four slabs per box / ghost planes: half the cases (run time)
            cases = cases[(op % 2)::2]
        for types, shape in cases:
            bad = emu.fe_one_level(op, nc, *shape, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 13 * nc + op)
            assert bad == 0, (types, shape)
        assert emu.fe_violations() == 0
    finally:
        emu.dh_set_planes(0)


@pytest.mark.parametrize("planes", [0, 1, 2])
@pytest.mark.parametrize("op,lam", [(1, 0.0), (3, 2.5)])
def test_ring_smoother_matches_general_path(emu, op, lam, planes):
    """k_smooth_ring, the persistent version of the half-sweep: 3 emulated CTAs of 8 consumer warps + 1 producer warp walk
    the tiles through a 5-stage ring with real mbarrier semantics (full / empty barriers, expect_tx byte counts, 8-byte
    cp.async arrivals); physical-boundary ghosts are built in place from the box's own first layer, plane faces read by
    the consumers.  Same bits as the general path, for whole levels (more tiles than stages), box lists and the
    in-kernel exchange between emulated ranks."""
    emu.fe_ranks.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    emu.dh_set_planes(planes)
    emu.fe_set_ring(2)
    try:
        for types, shape in zip(BC_SETS, [(2, 2, 1), (1, 2, 2), (5, 2, 2), (3, 1, 1)]):
            bad = emu.fe_one_level(op, 16, *shape, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 211 + op)
            assert bad == 0, (types, shape)
        if planes == 0:
            for types, shape, n_ranks in zip(BC_SETS[:3], [(2, 2, 2), (3, 2, 2), (4, 2, 3)], [2, 3, 4]):
                for bnd_first in (0, 1):
                    bad = emu.fe_ranks(op, 16, *shape, n_ranks, bnd_first, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 97 + op)
                    assert bad == 0, (types, shape, n_ranks, bnd_first)
        assert emu.fe_violations() == 0
    finally:
        emu.fe_set_ring(1)
        emu.dh_set_planes(0)


@pytest.mark.parametrize("planes", [0, 1, 2])
@pytest.mark.parametrize("op,lam", OPS)
@pytest.mark.parametrize("nc", [16, 32])
def test_fast_transfer_kernels_match_general_path(emu, nc, op, lam, planes):
    """update_coarse's residual + restriction and correct_children on a coarse level and the level of its octants"""
    emu.dh_set_planes(planes)
    try:
        for types, shape in zip(BC_SETS[2:], [(1, 1, 1), (2, 1, 1)]):
            bad = emu.fe_two_levels(op, nc, *shape, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 5 * nc + op)
            assert bad == 0, (types, shape)
        assert emu.fe_violations() == 0
    finally:
        emu.dh_set_planes(0)


@pytest.mark.parametrize("op,lam", OPS)
@pytest.mark.parametrize("nc", [16, 32])
def test_in_kernel_halo_exchange_between_emulated_ranks(emu, nc, op, lam):
    """the smoother's in-kernel exchange (PutDev: partition-face boxes store their fresh face layers into the neighbour
    ranks' double-buffered plane pools, the last such CTA signals and bumps the epoch) with the
    ranks' pools and flags as plain host pointers: four fused half-sweeps on 2, 3 and 4 emulated ranks follow the
    single-rank general path bit for bit, in box order and with the partition-face boxes launched first"""
    emu.fe_ranks.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    cases = list(zip(BC_SETS, [(2, 2, 2), (3, 2, 2), (2, 2, 1), (1, 2, 3)], [2, 3, 4, 2]))
    for types, shape, n_ranks in (cases if nc == 16 else cases[1:3]):   # box 32: four slabs per box, fewer cases
        for bnd_first in (0, 1):
            bad = emu.fe_ranks(op, nc, *shape, n_ranks, bnd_first, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 3 * nc + op)
            assert bad == 0, (types, shape, n_ranks, bnd_first)
    assert emu.fe_violations() == 0


@pytest.mark.parametrize("planes", [0, 1])
@pytest.mark.parametrize("op,lam", OPS + [(5, 1.25)])
def test_tail_interpreter_runs_a_vcycle_like_the_general_path(emu, op, lam, planes):
    """k_tail (omg_tail.cuh) with 1024 / 256 emulated threads: the V(2,2) op list of the coarse levels incl. the coarse solve and
    its residual test, on single-box and two-box hierarchies, general and fast (simple_*) per-cell functions, with and
    without staging the levels and their tables in shared memory -- state of every level, coarse iteration count and
    maximum residual equal to the ops applied one by one through the general functions"""
    emu.fe_tail.argtypes = [C.c_int] * 8 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    emu.dh_set_planes(planes)
    try:
        # (four- and three-box coarsest levels: the warp-level coarse solve takes up to four 2^3 boxes, three is the odd case
        # whose last eight lanes idle)
        cases = list(zip(BC_SETS + BC_SETS[2:], [(4, (1, 1, 1)), (4, (2, 1, 1)), (3, (1, 2, 1)), (3, (1, 1, 2)), (3, (2, 2, 1)), (3, (3, 1, 1))]))
        for types, (n_levels, shape) in (cases[2:] if planes else cases):
            for simple in (0, 1):
                for staged in (0, 1):
                    if staged and n_levels == 4 and shape != (1, 1, 1):
                        continue                      # four variables of all levels exceed the shared-memory budget: not staged
                    if shape in ((2, 2, 1), (3, 1, 1)) and (simple == 0 or staged == 0):
                        continue                      # these two exist for the warp-level coarse solve (fast path, as launched)
                    # 1024 threads as the product launches it for the staged single-box case, 256 elsewhere (run time)
                    n_threads = 1024 if (staged and shape == (1, 1, 1)) else 256
                    bad = emu.fe_tail(op, n_threads, n_levels, *shape, simple, staged, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 7 + op)
                    assert bad == 0, (types, n_levels, shape, simple, staged)
        assert emu.fe_violations() == 0
    finally:
        emu.dh_set_planes(0)


@pytest.mark.parametrize("planes", [0, 1, 2])
@pytest.mark.parametrize("op,lam", [(1, 0.0), (3, 2.5)])
@pytest.mark.parametrize("nc", [16, 8])
def test_gs_sweep_kernel_matches_general_path(emu, nc, op, lam, planes):
    """mg_smoother_gs: k_gs_sweep_fast (box staged in shared memory, hyperplanes i + j + k = const, one thread per (j, k)
    column) against a plain lexicographic sweep of gs_at over the same ghost snapshot: same bits after two sweeps."""
    emu.fe_gs_level.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
    emu.dh_set_planes(planes)
    try:
        for types, shape in zip(BC_SETS, [(2, 2, 1), (1, 2, 2), (2, 1, 2), (3, 1, 1)]):
            bad = emu.fe_gs_level(op, nc, *shape, six(C.c_int, types), six(C.c_double, BC_VALUES), lam, 97 * nc + op)
            assert bad == 0, (types, shape)
        assert emu.fe_violations() == 0
    finally:
        emu.dh_set_planes(0)
