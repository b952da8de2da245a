"""CPU tests of the host side: the C++ mirror of the reference's setup modules (tree, load balance)
against the oracle's independent restatement, and the C-ABI library surface.  No GPU needed: nothing
here launches a kernel (mg_allocate_storage must fail loudly without a device)."""
import ctypes
import os
import re

import numpy as np
import pytest

import problems as P
from compare import assert_trees_equal

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_uniform(cls, n, box, n_cpu=1, periodic=(False,) * 3, lb="default", smoother=P.mg_smoother_gsrb, **kw):
    mg = cls(P.mg_laplacian, smoother, n_cpu=n_cpu, **kw)
    mg.mg_set_methods()
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    mg.mg_build_rectangle(n3, box, [1.0 / n3[0]] * 3, [0.0] * 3, list(periodic))
    if lb == "simple":
        mg.mg_load_balance_simple()
    else:
        mg.mg_load_balance()
        if lb == "parents":
            mg.mg_load_balance_parents()
    return mg


@pytest.mark.parametrize("n,box", [(64, 16), (128, 16), (64, 32), ((64, 32, 32), 8), ((48, 48, 48), 8), ((32, 64, 96), 16), (16, 2)])
@pytest.mark.parametrize("n_cpu", [1, 2, 5, 8])
def test_rectangle_tree_and_partition_match_oracle(oracle, product, n, box, n_cpu):
    for lb in ("default", "simple", "parents"):
        a = build_uniform(oracle.OracleMg, n, box, n_cpu=n_cpu, lb=lb)
        b = build_uniform(product.Mg, n, box, n_cpu=n_cpu, lb=lb)
        assert_trees_equal(a, b)


def test_periodic_and_gs_trees(oracle, product):
    for per in [(True, True, True), (True, False, True)]:
        assert_trees_equal(build_uniform(oracle.OracleMg, 32, 8, periodic=per), build_uniform(product.Mg, 32, 8, periodic=per))
    a = build_uniform(oracle.OracleMg, 64, 16, smoother=P.mg_smoother_gs)
    b = build_uniform(product.Mg, 64, 16, smoother=P.mg_smoother_gs)
    assert_trees_equal(a, b)
    assert a.box_size_lvl(a.lowest_lvl) == 2


@pytest.mark.parametrize("n_cpu", [1, 3, 8])
def test_amr_tree_matches_oracle(oracle, product, n_cpu):
    def amr(cls, **kw):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb, n_cpu=n_cpu, **kw)
        mg.mg_set_methods()
        mg.mg_allocate_storage = lambda: None         # tree only (no device in this test)
        P.build_amr_tree(mg, 3, [64] * 3, 16, [1.0 / 64] * 3)
        return mg
    a, b = amr(oracle.OracleMg), amr(product.Mg)
    assert_trees_equal(a, b)
    assert [len(b.ids(l)) for l in (1, 2, 3)] == [64, 64, 64]
    assert len(b.ref_bnds(1)) > 0 and len(b.ref_bnds(2)) > 0


def test_partition_is_morton_contiguous(product):
    """north-star: boxes partitioned in Morton order as in m_load_balance -- 8 ranks = 8 octants"""
    mg = build_uniform(product.Mg, 128, 16, n_cpu=8)
    t = mg.tree_export()
    ids = mg.ids(1)
    ranks = [t[i - 1, 0] for i in ids]
    assert ranks == sorted(ranks)                                     # contiguous chunks of the ids list
    for i in ids:
        ix = t[i - 1, 2:5]
        octant = (ix[0] - 1) // 4 + 2 * ((ix[1] - 1) // 4) + 4 * ((ix[2] - 1) // 4)
        assert t[i - 1, 0] == octant
    assert mg.mg_number_of_unknowns() == 128 ** 3


def test_no_device_fails_loudly(product):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    mg = build_uniform(product.Mg, 32, 8)
    with pytest.raises(product.MgError, match="no CUDA device"):
        mg.mg_allocate_storage()


def test_c_abi_exports_every_declared_symbol(product):
    """every function declared in include/omg_gpu.h is exported by libomg_b200.so"""
    hdr = open(os.path.join(ROOT, "include", "omg_gpu.h")).read()
    names = sorted(set(re.findall(r"\b(omg_[a-z_0-9]+)\s*\(", hdr)))
    assert len(names) >= 25
    G, _ = product.libs()
    for n in names:
        assert hasattr(G, n), n
    assert b"sm_100a" in G.omg_version()
    h = ctypes.c_void_p()
    import torch
    if not torch.cuda.is_available():
        assert G.omg_create(0, ctypes.byref(h)) != 0 and not h.value    # no CPU fallback


def test_product_does_not_reference_the_oracle():
    """the product tree must not import / link / mention the oracle"""
    pkg = os.path.join(ROOT, "octree-mg_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.path.basename(dirpath) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", ".f90")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in src and "import oracle" not in src and "from oracle" not in src, f


def _build_cpp_example(name="test_uniform_grid"):
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "examples", name)
    subprocess.check_call(["g++", "-O2", "-std=c++17", os.path.join(root, "examples", name + ".cpp"), "-o", exe,
                           "-L" + os.path.join(root, "octree-mg_b200"), "-lomg_host", "-lomg_b200",
                           "-Wl,-rpath," + os.path.join(root, "octree-mg_b200")])
    return exe


def test_cpp_host_example_builds_and_fails_loudly_without_a_device(product):
    """examples/test_uniform_grid.cpp = the reference's tests/test_uniform_grid.f90 against the C++ mirror: it must
    compile and link against the two libraries, and on a machine without a GPU stop like the reference's `error stop`"""
    import subprocess
    import torch
    exe = _build_cpp_example()
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: the example would run (see the gpu-marked test)")
    r = subprocess.run([exe, "16", "32", "32", "32", "2", "t", "gsrb"], capture_output=True, text=True)
    assert r.returncode == 1
    assert "no CUDA device available (there is no CPU fallback)" in r.stderr


def test_cpp_performance_example_builds_and_fails_loudly_without_a_device(product):
    """examples/test_performance.cpp = the reference's tests/test_performance.f90 (the program of the headline number)"""
    import subprocess
    import torch
    exe = _build_cpp_example("test_performance")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "Usage" in r.stderr
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: the example would run (see the gpu-marked test)")
    r = subprocess.run([exe, "16", "32", "32", "32", "2", "f", "gsrb"], capture_output=True, text=True)
    assert r.returncode == 1
    assert "no CUDA device available (there is no CPU fallback)" in r.stderr


def test_cpp_refinement_example_builds_and_fails_loudly_without_a_device(product):
    """examples/test_refinement.cpp = the reference's tests/test_refinement.f90: the AMR tree is built through the mirror's
    mg_add_children / mg_set_* calls before storage is allocated, so the tree code runs here even without a GPU"""
    import subprocess
    import torch
    exe = _build_cpp_example("test_refinement")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "Usage" in r.stderr
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: the example would run (see the gpu-marked test)")
    r = subprocess.run([exe, "3", "8", "32", "32", "32", "2", "t", "gsrb"], capture_output=True, text=True)
    assert r.returncode == 1
    assert "no CUDA device available (there is no CPU fallback)" in r.stderr


@pytest.mark.gpu
def test_cpp_refinement_example_converges_on_the_device():
    import re
    import subprocess
    exe = _build_cpp_example("test_refinement")
    r = subprocess.run([exe, "3", "8", "32", "32", "32", "8", "t", "gsrb"], capture_output=True, text=True, check=True)
    errs = [float(m.group(1)) for m in re.finditer(r"max err\s+(\S+)", r.stdout)]
    # the oracle's history for this set-up: 0.999, 1.0e-2, 2.3e-4, 1.3e-5, 8.7e-7, 6.0e-8, 4.2e-9, 2.9e-10, 2.0e-11
    assert len(errs) == 9 and 0.9 < errs[0] < 1.0 and errs[-1] < 1e-9, r.stdout
    assert "lvl_3: 64 boxes, 64 leaves, box size: 8" in r.stdout, r.stdout


@pytest.mark.gpu
def test_cpp_performance_example_runs_on_the_device():
    import re
    import subprocess
    exe = _build_cpp_example("test_performance")
    for fmg, sm in (("f", "gsrb"), ("t", "gs")):
        r = subprocess.run([exe, "16", "64", "64", "64", "5", fmg, sm], capture_output=True, text=True, check=True)
        assert re.search(r"unknowns/microsec\s+\S+", r.stdout), r.stdout
        res = float(re.search(r"max residual\s+(\S+)", r.stdout).group(1))
        assert 0.0 <= res < 1e-3, r.stdout            # rhs = 1: five cycles take the residual far below its initial 1


@pytest.mark.gpu
def test_cpp_host_example_converges_on_the_device():
    import re
    import subprocess
    exe = _build_cpp_example()
    r = subprocess.run([exe, "16", "64", "64", "64", "6", "t", "gsrb"], capture_output=True, text=True, check=True)
    errs = [float(m.group(1)) for m in re.finditer(r"max solution error\s+(\S+)", r.stdout)]
    assert len(errs) == 7 and errs[-1] < 1e-9 * max(1.0, errs[0])      # the reference test's pass criterion (:132-202)


def test_host_side_error_stops_keep_the_reference_wording(product):
    """SURVEY 8b "error convention": the reference `error stop`s with a text; the mirror raises with the same text
    (src/m_build_tree.f90:32-34, src/m_multigrid.f90:50, src/m_helmholtz.f90:42-43, src/m_allocate_storage.f90:63,
    src/m_laplacian.f90:46).  All of these are decided on the host: no device needed."""
    def fresh(op=P.mg_laplacian):
        mg = product.Mg(op, P.mg_smoother_gsrb)
        for nb in range(1, 7):
            mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.0)
        return mg

    with pytest.raises(product.MgError, match="box_size should be even"):
        fresh().mg_build_rectangle([30, 30, 30], 5, [1 / 30.0] * 3, [0.0] * 3, [False] * 3)
    with pytest.raises(product.MgError, match="box_size does not divide domain_size"):
        fresh().mg_build_rectangle([32, 32, 20], 8, [1 / 32.0] * 3, [0.0] * 3, [False] * 3)
    with pytest.raises(product.MgError, match="mg_set_methods: unknown operator"):
        fresh(op=9).mg_set_methods()
    with pytest.raises(product.MgError, match="lambda < 0 not allowed"):
        fresh(op=P.mg_helmholtz).set_lambda(-1.0)
    with pytest.raises(product.MgError, match="allocate_storage: tree is not yet created"):
        fresh().mg_allocate_storage()
    mg = fresh()
    mg.set_int("geometry_type", 2)                      # mg_cylindrical: 2-D only in the reference, unsupported here
    with pytest.raises(product.MgError, match="unsupported geometry"):
        mg.mg_set_methods()
    mg = fresh()
    mg.set_int("smoother_type", 3)                      # mg_smoother_jacobi is declared but not implemented by the reference either
    with pytest.raises(product.MgError, match="unsupported smoother type"):
        mg.mg_set_methods()


def check_tree_against_geometry(mg, n3, periodic):
    """Pins the tree to plain geometry, without the libraries' index bookkeeping (ix, child offsets, id arithmetic): every
    box is a cube [r_min, r_min + nc * dr); its neighbour through a face is whichever box of the level contains the point
    one box size beyond its centre (the periodic image of that point; outside the domain: a physical boundary, id < 0;
    inside but not covered: a refinement boundary, id 0 = mg_no_box, src/m_data_structures.f90:60-66); its parent is the box
    of the level below that contains its centre, its children those of the level above whose centres it contains, x fastest
    (src/m_build_tree.f90:330-393, mg_child_dix src/m_data_structures.f90:150-170)."""
    L = np.array(n3, dtype=float) * mg.dr(1)                      # domain size (r_min = 0)
    boxes = {}
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        size = mg.box_size_lvl(lvl) * mg.dr(lvl)
        assert len(set(mg.ids(lvl))) == len(mg.ids(lvl))
        for id_ in mg.ids(lvl):
            info = mg.box_info(id_)
            assert info["lvl"] == lvl
            key = tuple(np.rint(info["r_min"] / size).astype(int))
            assert np.array_equal(np.array(key) * size, info["r_min"]), "boxes are aligned to their own size"
            assert (lvl, key) not in boxes, "two boxes of a level overlap"
            boxes[(lvl, key)] = (id_, info, size)
    per_level = {}
    for (lvl, key), v in boxes.items():
        per_level.setdefault(lvl, {})[key] = v

    def find(lvl, point):
        if lvl not in per_level:
            return None
        size = next(iter(per_level[lvl].values()))[2]
        key = tuple(np.floor(point / size).astype(int))
        hit = per_level[lvl].get(key)
        return hit[0] if hit else None

    n_checked = 0
    for (lvl, key), (id_, info, size) in boxes.items():
        centre = info["r_min"] + 0.5 * size
        for nb in range(6):
            d, sign = nb // 2, (1 if nb % 2 else -1)
            p = centre.copy()
            p[d] += sign * size[d]
            got = info["neighbors"][nb]
            if p[d] < 0 or p[d] > L[d]:
                if periodic[d]:
                    p[d] = p[d] % L[d]
                    assert got == find(lvl, p), (lvl, key, nb)
                else:
                    assert got < 0, (lvl, key, nb, got)
            else:
                want = find(lvl, p)
                assert got == (want if want is not None else 0), (lvl, key, nb, got, want)
            n_checked += 1
        want_parent = find(lvl - 1, centre)
        if lvl > mg.lowest_lvl:
            assert info["parent"] == want_parent, (lvl, key)
        kids = [c for c in info["children"] if c > 0]
        if lvl < mg.highest_lvl and lvl + 1 in per_level:
            fsize = next(iter(per_level[lvl + 1].values()))[2]
            if np.allclose(fsize, size):          # below first_normal_lvl: one child, the same cube at twice the resolution
                want_kids = [find(lvl + 1, centre)]
            else:
                want_kids = [find(lvl + 1, info["r_min"] + (np.array([cx, cy, cz]) + 0.5) * fsize)
                             for cz in (0, 1) for cy in (0, 1) for cx in (0, 1)]
            want_kids = [k for k in want_kids if k is not None]
            assert kids == want_kids, (lvl, key, kids, want_kids)
        else:
            assert kids == []
    return n_checked


@pytest.mark.parametrize("n,box,periodic", [(64, 16, (False,) * 3), ((64, 32, 32), 8, (False,) * 3), ((32, 64, 96), 16, (False, True, False)),
                                            (32, 8, (True,) * 3), ((48, 48, 48), 8, (False,) * 3)])
def test_uniform_tree_matches_plain_geometry(oracle, product, n, box, periodic):
    for cls in (oracle.OracleMg, product.Mg):
        mg = build_uniform(cls, n, box, periodic=periodic)
        n3 = [n] * 3 if np.isscalar(n) else list(n)
        assert check_tree_against_geometry(mg, n3, periodic) >= 6 * len(mg.ids(1))


def test_refined_tree_matches_plain_geometry(oracle, product):
    for cls in (oracle.OracleMg, product.Mg):
        mg = cls(P.mg_laplacian, P.mg_smoother_gsrb)
        mg.mg_set_methods()
        mg.mg_allocate_storage = lambda: None         # tree only
        P.build_amr_tree(mg, 3, [64] * 3, 16, [1.0 / 64] * 3)
        check_tree_against_geometry(mg, [64] * 3, (False,) * 3)
        # refinement boundaries: exactly the faces of the refined cubes that do not touch another fine box
        n_rb = sum(1 for lvl in (2, 3) for id_ in mg.ids(lvl) for nbr in mg.box_info(id_)["neighbors"] if nbr == 0)
        assert n_rb == 2 * 6 * 16                     # two refined cubes of 4^3 boxes: 6 faces x 4 x 4 box faces each
