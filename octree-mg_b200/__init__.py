"""octree-mg_b200: B200-native (sm_100a) implementation of octree-mg's multigrid-cycle hot path.

Python is harness plumbing only: this module loads the native libraries and exposes the C++ host
mirror of the reference's Fortran API (`Mg`, same routine names as src/m_octree_mg.f90) for the
tests and the benchmark.  The product is libomg_b200.so (CUDA kernels behind include/omg_gpu.h);
nothing here computes on the CPU and nothing here imports the oracle.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# constants of src/m_data_structures.f90:8-87
mg_laplacian, mg_vlaplacian, mg_helmholtz, mg_vhelmholtz, mg_ahelmholtz = 1, 2, 3, 4, 5
mg_smoother_gs, mg_smoother_gsrb = 1, 2
mg_iphi, mg_irhs, mg_iold, mg_ires, mg_iveps = 1, 2, 3, 4, 5
mg_num_vars = 4
mg_bc_dirichlet, mg_bc_neumann, mg_bc_continuous = -10, -11, -12
mg_num_neighbors = 6

BC_CALLBACK = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int,
                          C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_void_p)

_libs = None


class MgError(RuntimeError):
    pass


class omg_tree_desc(C.Structure):
    _fields_ = [("n_boxes", C.c_int32), ("box_size", C.c_int32), ("lowest_lvl", C.c_int32), ("highest_lvl", C.c_int32),
                ("first_normal_lvl", C.c_int32), ("my_rank", C.c_int32), ("n_cpu", C.c_int32),
                ("n_extra_vars", C.c_int32), ("periodic", C.c_int32 * 3),
                ("box_size_lvl", C.POINTER(C.c_int32)), ("dr", C.POINTER(C.c_double)),
                ("box_rank", C.POINTER(C.c_int32)), ("box_lvl", C.POINTER(C.c_int32)), ("box_ix", C.POINTER(C.c_int32)),
                ("box_parent", C.POINTER(C.c_int32)), ("box_children", C.POINTER(C.c_int32)),
                ("box_neighbors", C.POINTER(C.c_int32)), ("lvl_ptr", C.POINTER(C.c_int32)),
                ("lvl_ids", C.POINTER(C.c_int32))]


class omg_params(C.Structure):
    _fields_ = [("operator_type", C.c_int32), ("smoother_type", C.c_int32), ("geometry_type", C.c_int32),
                ("n_cycle_down", C.c_int32), ("n_cycle_up", C.c_int32), ("max_coarse_cycles", C.c_int32),
                ("subtract_mean", C.c_int32), ("residual_coarse_abs", C.c_double), ("residual_coarse_rel", C.c_double),
                ("lambda_", C.c_double)]


def build(force=False):
    """Compile libomg_b200.so / libomg_host.so in-tree (nvcc, sm_100a)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_omg_build", os.path.join(HERE, "build.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.build(force=force)


def libs():
    """(libomg_b200, libomg_host) ctypes handles.  Fails loudly if the native code is not built."""
    global _libs
    if _libs is None:
        p_cuda = os.path.join(HERE, "libomg_b200.so")
        p_host = os.path.join(HERE, "libomg_host.so")
        if not (os.path.exists(p_cuda) and os.path.exists(p_host)):
            raise MgError("native libraries missing: run `python octree-mg_b200/build.py` (there is no CPU fallback)")
        # one NCCL per process: bind the copy PyTorch ships (if any) so that both can live together
        nccl = os.path.join(os.path.dirname(os.path.dirname(np.__file__)), "nvidia", "nccl", "lib", "libnccl.so.2")
        if os.path.exists(nccl):
            os.environ.setdefault("OMG_NCCL_LIB", nccl)
        G = C.CDLL(p_cuda, mode=C.RTLD_GLOBAL)
        Hh = C.CDLL(p_host)
        _declare(G, Hh)
        _libs = (G, Hh)
    return _libs


def _declare(G, H):
    vp, ci, cd = C.c_void_p, C.c_int, C.c_double
    G.omg_create.argtypes = [ci, C.POINTER(vp)]
    G.omg_destroy.argtypes = [vp]
    G.omg_last_error.restype = C.c_char_p
    G.omg_last_error.argtypes = [vp]
    G.omg_version.restype = C.c_char_p
    G.omg_set_tree.argtypes = [vp, C.POINTER(omg_tree_desc)]
    G.omg_set_params.argtypes = [vp, C.POINTER(omg_params)]
    G.omg_set_bc_const.argtypes = [vp, ci, ci, ci, cd]
    G.omg_set_bc_plane.argtypes = [vp, ci, ci, ci, C.POINTER(cd)]
    G.omg_push_var.argtypes = [vp, ci, ci, C.POINTER(C.c_int32), C.POINTER(C.POINTER(cd))]
    G.omg_pull_var.argtypes = [vp, ci, ci, C.POINTER(C.c_int32), C.POINTER(C.POINTER(cd)), ci]
    G.omg_push_var_packed.argtypes = [vp, ci, ci, ci, vp, C.c_int64]
    G.omg_pull_var_packed.argtypes = [vp, ci, ci, ci, vp, C.c_int64, ci]
    G.omg_packed_size.restype = C.c_int64
    G.omg_packed_size.argtypes = [vp, ci, ci]
    G.omg_set_var_const.argtypes = [vp, ci, ci, ci, cd, ci]
    G.omg_copy_to_tree.argtypes = [vp, ci, ci, ci, ci, ci, vp, C.c_int64]
    G.omg_copy_from_tree.argtypes = [vp, ci, ci, ci, ci, ci, vp, C.c_int64]
    G.omg_copy_size.restype = C.c_int64
    G.omg_copy_size.argtypes = [vp, ci, ci, ci, ci]
    G.omg_fas_vcycle.argtypes = [vp, ci, ci, ci, ci, C.POINTER(cd)]
    G.omg_fas_fmg.argtypes = [vp, ci, ci, C.POINTER(cd)]
    G.omg_fas_vcycles.argtypes = [vp, ci, C.POINTER(cd)]
    G.omg_fill_ghost_cells_lvl.argtypes = [vp, ci, ci]
    G.omg_debug_step.argtypes = [vp, C.c_char_p, ci, ci, C.POINTER(cd)]
    G.omg_restrict_lvl.argtypes = [vp, ci, ci]
    G.omg_prolong.argtypes = [vp, ci, ci, ci, ci]
    G.omg_apply_op.argtypes = [vp, ci]
    G.omg_diffusion_set_rhs.argtypes = [vp, cd, cd]
    G.omg_debug_ddiv.argtypes = [ci, vp, vp, vp, C.c_int64, ci]
    G.omg_debug_ddiv_static.argtypes = [ci, vp, vp, vp, C.c_int64, ci, C.c_char_p, ci]
    G.omg_nccl_unique_id.argtypes = [C.c_char_p]
    G.omg_comm_init.argtypes = [vp, C.c_char_p, ci, ci]
    G.omg_comm_init_local.argtypes = [C.POINTER(vp), ci]
    G.omg_exchange_plan.argtypes = [C.POINTER(omg_tree_desc), ci, ci, ci, ci, C.POINTER(C.c_int64), C.POINTER(C.c_int64), ci,
                                    C.POINTER(ci), C.POINTER(ci)]
    G.omg_sync.argtypes = [vp]
    G.omg_stat.restype = C.c_int64
    G.omg_stat.argtypes = [vp, C.c_char_p]
    G.omg_set_option.argtypes = [vp, C.c_char_p, ci]
    G.omg_trace.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    G.omg_step_name.restype = C.c_char_p
    G.omg_step_name.argtypes = [C.c_int]
    G.omg_last_ms.restype = cd
    G.omg_last_ms.argtypes = [vp, C.c_char_p]
    G.omg_stream.restype = vp
    G.omg_stream.argtypes = [vp]
    G.omg_timers.argtypes = [vp, C.POINTER(ci), C.POINTER(C.c_char_p), C.POINTER(cd)]

    H.omgh_create.restype = vp
    H.omgh_create.argtypes = [ci]
    H.omgh_destroy.argtypes = [vp]
    H.omgh_last_error.restype = C.c_char_p
    H.omgh_last_error.argtypes = [vp]
    H.omgh_gpu_handle.restype = vp
    H.omgh_gpu_handle.argtypes = [vp]
    H.omgh_set_int.argtypes = [vp, C.c_char_p, ci]
    H.omgh_get_int.argtypes = [vp, C.c_char_p]
    H.omgh_set_real.argtypes = [vp, C.c_char_p, cd]
    H.omgh_number_of_unknowns.restype = C.c_longlong
    H.omgh_number_of_unknowns.argtypes = [vp]
    H.omgh_box_size_lvl.argtypes = [vp, ci]
    H.omgh_dr_lvl.argtypes = [vp, ci, C.POINTER(cd)]
    H.omgh_set_bc.argtypes = [vp, ci, ci, ci, cd]
    H.omgh_set_bc_callback.argtypes = [vp, ci, ci, BC_CALLBACK, vp]
    H.omgh_comm_init.argtypes = [vp, ci, ci]
    H.omgh_build_rectangle.argtypes = [vp, C.POINTER(ci), ci, C.POINTER(cd), C.POINTER(cd), C.POINTER(ci), ci]
    for n in ("omgh_set_methods", "omgh_allocate_storage", "omgh_deallocate_storage", "omgh_gpu_update_tree", "omgh_phi_bc_store",
              "omgh_gpu_push_bc", "omgh_gpu_push_params"):
        getattr(H, n).argtypes = [vp]
    for n in ("omgh_add_children", "omgh_set_leaves_parents", "omgh_set_next_level_ids", "omgh_set_neighbors_lvl",
              "omgh_set_refinement_boundaries", "omgh_load_balance", "omgh_apply_op", "omgh_fill_ghost_cells",
              "omgh_restrict", "omgh_gpu_push"):
        getattr(H, n).argtypes = [vp, ci]
    for n in ("omgh_fill_ghost_cells_lvl", "omgh_restrict_lvl", "omgh_set_box_rank", "omgh_gpu_pull"):
        getattr(H, n).argtypes = [vp, ci, ci]
    H.omgh_prolong.argtypes = [vp, ci, ci, ci, ci]
    H.omgh_lvl_list.argtypes = [vp, ci, ci, C.POINTER(ci)]
    H.omgh_tree_export.argtypes = [vp, C.POINTER(ci)]
    H.omgh_box_rmin.argtypes = [vp, ci, C.POINTER(cd)]
    H.omgh_box_cc.restype = C.POINTER(cd)
    H.omgh_box_cc.argtypes = [vp, ci]
    H.omgh_fas_vcycle.argtypes = [vp, ci, ci, ci, C.POINTER(cd), ci]
    H.omgh_fas_fmg.argtypes = [vp, ci, ci, C.POINTER(cd)]
    H.omgh_diffusion_solve.argtypes = [vp, ci, cd, cd, ci, cd, C.POINTER(cd), C.POINTER(ci)]
    H.omgh_poisson_free_3d.argtypes = [vp, ci, cd, ci, ci, C.POINTER(cd)]
    H.omgh_free_space_fft_lvl.argtypes = [vp]


def _arr(vals, ctype):
    return (ctype * len(vals))(*vals)


class Mg:
    """mg_t of the reference, backed by the C++ host mirror and the GPU library.

    gpu_resident=False (default) is the strict drop-in mode: every mg_fas_* call pushes phi/rhs and
    pulls phi/rhs/old/res, so `cc()` views behave like mg%boxes(id)%cc in the reference.  With
    gpu_resident=True the caller brackets its own cc accesses with mg_gpu_push / mg_gpu_pull.
    """

    def __init__(self, operator_type=mg_laplacian, smoother_type=mg_smoother_gsrb, n_extra_vars=0, n_cpu=1, my_rank=0,
                 device=0, gpu_resident=False):
        self.G, self.H = libs()
        self.h = C.c_void_p(self.H.omgh_create(device))
        self._keep = []
        self._tree = None
        self.set_int("operator_type", operator_type)
        self.set_int("smoother_type", smoother_type)
        self.set_int("n_extra_vars", n_extra_vars)
        self.set_int("gpu_resident", int(gpu_resident))
        self.H.omgh_comm_init(self.h, my_rank, n_cpu)

    def __del__(self):
        try:
            self.H.omgh_destroy(self.h)
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise MgError(self.H.omgh_last_error(self.h).decode())

    @property
    def gpu(self):
        """omg_handle* of the C ABI (None before mg_allocate_storage)."""
        p = self.H.omgh_gpu_handle(self.h)
        return C.c_void_p(p) if p else None

    # -- parameters -----------------------------------------------------------------------------
    def set_int(self, key, v):
        self._ck(self.H.omgh_set_int(self.h, key.encode(), int(v)))

    def get_int(self, key):
        return self.H.omgh_get_int(self.h, key.encode())

    def set_real(self, key, v):
        self._ck(self.H.omgh_set_real(self.h, key.encode(), float(v)))

    def set_lambda(self, lam):
        self.set_real("lambda", lam)

    def set_bc(self, nb, iv, bc_type, bc_value=0.0):
        self._ck(self.H.omgh_set_bc(self.h, nb, iv, bc_type, bc_value))

    def set_bc_callback(self, nb, iv, fn):
        """fn(r_min[3], dr[3], lvl, nc, iv, nb) -> (bc_type, bc[nc, nc] indexed [t2, t1])."""
        def tramp(r_min, dr, lvl, nc, iv_, nb_, bc_type, bc, user):
            t, plane = fn(np.array(r_min[0:3]), np.array(dr[0:3]), lvl, nc, iv_, nb_)
            bc_type[0] = t
            np.ctypeslib.as_array(bc, shape=(nc, nc))[...] = plane
        cb = BC_CALLBACK(tramp)
        self._keep.append(cb)
        self._ck(self.H.omgh_set_bc_callback(self.h, nb, iv, cb, None))

    # -- tree / setup -----------------------------------------------------------------------------
    def mg_set_methods(self):
        self._ck(self.H.omgh_set_methods(self.h))

    def mg_build_rectangle(self, domain_size, box_size, dx, r_min, periodic, n_finer=0):
        self._ck(self.H.omgh_build_rectangle(self.h, _arr(list(domain_size), C.c_int), box_size,
                                             _arr(list(dx), C.c_double), _arr(list(r_min), C.c_double),
                                             _arr([int(bool(p)) for p in periodic], C.c_int), n_finer))
        self._tree = None

    def mg_add_children(self, id_):
        self._ck(self.H.omgh_add_children(self.h, id_))
        self._tree = None

    def mg_set_leaves_parents(self, lvl):
        self._ck(self.H.omgh_set_leaves_parents(self.h, lvl))

    def mg_set_next_level_ids(self, lvl):
        self._ck(self.H.omgh_set_next_level_ids(self.h, lvl))

    def mg_set_neighbors_lvl(self, lvl):
        self._ck(self.H.omgh_set_neighbors_lvl(self.h, lvl))
        self._tree = None

    def mg_set_refinement_boundaries(self, lvl):
        self._ck(self.H.omgh_set_refinement_boundaries(self.h, lvl))

    def mg_load_balance_simple(self):
        self._ck(self.H.omgh_load_balance(self.h, 0))
        self._tree = None

    def mg_load_balance(self):
        self._ck(self.H.omgh_load_balance(self.h, 1))
        self._tree = None

    def mg_load_balance_parents(self):
        self._ck(self.H.omgh_load_balance(self.h, 2))
        self._tree = None

    def mg_allocate_storage(self):
        self._ck(self.H.omgh_allocate_storage(self.h))

    def mg_deallocate_storage(self):
        self._ck(self.H.omgh_deallocate_storage(self.h))

    def mg_gpu_update_tree(self):
        """the host changed the tree of an allocated mg (mg_add_children ..., mg_load_balance): on-device regrid"""
        self._ck(self.H.omgh_gpu_update_tree(self.h))
        self._tree = None

    # -- introspection ------------------------------------------------------------------------------
    lowest_lvl = property(lambda self: self.get_int("lowest_lvl"))
    first_normal_lvl = property(lambda self: self.get_int("first_normal_lvl"))
    n_boxes = property(lambda self: self.get_int("n_boxes"))
    box_size = property(lambda self: self.get_int("box_size"))
    n_extra_vars = property(lambda self: self.get_int("n_extra_vars"))

    @property
    def highest_lvl(self):
        return self.get_int("highest_lvl")

    @highest_lvl.setter
    def highest_lvl(self, v):
        self.set_int("highest_lvl", v)

    def mg_number_of_unknowns(self):
        return int(self.H.omgh_number_of_unknowns(self.h))

    def box_size_lvl(self, lvl):
        return self.H.omgh_box_size_lvl(self.h, lvl)

    def dr(self, lvl):
        out = (C.c_double * 3)()
        self.H.omgh_dr_lvl(self.h, lvl, out)
        return np.array(out[:])

    def _lst(self, lvl, which):
        n = self.H.omgh_lvl_list(self.h, lvl, which, None)
        out = (C.c_int * max(n, 1))()
        self.H.omgh_lvl_list(self.h, lvl, which, out)
        return list(out[:n])

    def ids(self, lvl):
        return self._lst(lvl, 0)

    def leaves(self, lvl):
        return self._lst(lvl, 1)

    def parents(self, lvl):
        return self._lst(lvl, 2)

    def ref_bnds(self, lvl):
        return self._lst(lvl, 3)

    def my_ids(self, lvl):
        return self._lst(lvl, 4)

    def my_leaves(self, lvl):
        return self._lst(lvl, 5)

    def tree_export(self):
        """(n_boxes, 20) int32: rank, lvl, ix(3), parent, children(8), neighbors(6) for ids 1..n_boxes."""
        out = np.zeros((self.n_boxes, 20), dtype=np.int32)
        self.H.omgh_tree_export(self.h, out.ctypes.data_as(C.POINTER(C.c_int)))
        return out

    def tree_desc(self):
        """(omg_tree_desc, keepalive): the flattened tree mg_allocate_storage hands to omg_set_tree; needs no device."""
        t = np.ascontiguousarray(self.tree_export())
        n = t.shape[0]
        cols = dict(rank=t[:, 0], lvl=t[:, 1], ix=t[:, 2:5], parent=t[:, 5], children=t[:, 6:14], neighbors=t[:, 14:20])
        keep = {k: np.ascontiguousarray(v, dtype=np.int32) for k, v in cols.items()}
        ptr, ids = [0] * 42, []
        for lvl in range(-20, 21):
            if self.lowest_lvl <= lvl <= self.highest_lvl:
                ids += self.ids(lvl)
            ptr[lvl + 20 + 1] = len(ids)
        keep["lvl_ptr"] = np.array(ptr, dtype=np.int32)
        keep["lvl_ids"] = np.array(ids, dtype=np.int32)
        keep["bsl"] = np.array([self.box_size_lvl(l) for l in range(-20, 21)], dtype=np.int32)
        keep["dr"] = np.array([self.dr(l) for l in range(-20, 21)], dtype=np.float64)
        p32 = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        d = omg_tree_desc(n, self.box_size, self.lowest_lvl, self.highest_lvl, self.first_normal_lvl,
                          self.get_int("my_rank"), self.get_int("n_cpu"), self.n_extra_vars, (C.c_int32 * 3)(0, 0, 0),
                          p32(keep["bsl"]), keep["dr"].ctypes.data_as(C.POINTER(C.c_double)), p32(keep["rank"]),
                          p32(keep["lvl"]), p32(keep["ix"]), p32(keep["parent"]), p32(keep["children"]),
                          p32(keep["neighbors"]), p32(keep["lvl_ptr"]), p32(keep["lvl_ids"]))
        return d, keep

    def exchange_plan(self, phase, lvl, peer, me=None):
        """(send_keys, recv_keys) of the static exchange plan (omg_exchange_plan); phase 0 ghost, 1 restrict, 2 prolong."""
        d, keep = self.tree_desc()
        me = self.get_int("my_rank") if me is None else me
        ns, nr = C.c_int(0), C.c_int(0)
        self.G.omg_exchange_plan(C.byref(d), me, phase, lvl, peer, None, None, 0, C.byref(ns), C.byref(nr))
        sk, rk = (C.c_int64 * max(ns.value, 1))(), (C.c_int64 * max(nr.value, 1))()
        self.G.omg_exchange_plan(C.byref(d), me, phase, lvl, peer, sk, rk, max(ns.value, nr.value, 1), C.byref(ns), C.byref(nr))
        return list(sk[:ns.value]), list(rk[:nr.value])

    def box_info(self, id_):
        if self._tree is None:
            self._tree = self.tree_export()
        o = [int(x) for x in self._tree[id_ - 1]]
        rmin = (C.c_double * 3)()
        self.H.omgh_box_rmin(self.h, id_, rmin)
        return dict(rank=o[0], lvl=o[1], ix=o[2:5], parent=o[5], children=o[6:14], neighbors=o[14:20],
                    r_min=np.array(rmin[:]))

    def set_box_rank(self, id_, rank):
        self.H.omgh_set_box_rank(self.h, id_, rank)
        self._tree = None

    def cc(self, id_, nc=None):
        """numpy view [v, k, j, i] of the host copy boxes(id)%cc (ghost cells included)."""
        if nc is None:
            nc = self.box_size_lvl(self.box_info(id_)["lvl"])
        nv = mg_num_vars + self.n_extra_vars
        p = self.H.omgh_box_cc(self.h, id_)
        if not p:
            raise MgError("box %d is not owned by this rank" % id_)
        return np.ctypeslib.as_array(p, shape=(nv, nc + 2, nc + 2, nc + 2))

    # -- hot path (device) ----------------------------------------------------------------------------
    def mg_fill_ghost_cells_lvl(self, lvl, iv):
        self._ck(self.H.omgh_fill_ghost_cells_lvl(self.h, lvl, iv))

    def mg_fill_ghost_cells(self, iv):
        self._ck(self.H.omgh_fill_ghost_cells(self.h, iv))

    def mg_restrict_lvl(self, iv, lvl):
        self._ck(self.H.omgh_restrict_lvl(self.h, iv, lvl))

    def mg_restrict(self, iv):
        self._ck(self.H.omgh_restrict(self.h, iv))

    def mg_prolong(self, lvl, iv, iv_to, add):
        self._ck(self.H.omgh_prolong(self.h, lvl, iv, iv_to, int(bool(add))))

    def mg_apply_op(self, i_out):
        self._ck(self.H.omgh_apply_op(self.h, i_out))

    def mg_phi_bc_store(self):
        self._ck(self.H.omgh_phi_bc_store(self.h))

    def mg_fas_vcycle(self, highest_lvl=None, want_res=True, standalone=True):
        r = C.c_double(0.0)
        has = highest_lvl is not None
        self._ck(self.H.omgh_fas_vcycle(self.h, int(has), highest_lvl if has else 0, int(want_res), C.byref(r),
                                        int(standalone)))
        return r.value if want_res else None

    def mg_fas_fmg(self, have_guess, want_res=True):
        r = C.c_double(0.0)
        self._ck(self.H.omgh_fas_fmg(self.h, int(bool(have_guess)), int(want_res), C.byref(r)))
        return r.value if want_res else None

    # src/m_diffusion.f90:19-157; return (last max residual, V-cycles added after the FMG cycle)
    def _diffusion(self, kind, dt, coeff, order, max_res):
        r, n = C.c_double(0.0), C.c_int(0)
        self._ck(self.H.omgh_diffusion_solve(self.h, kind, dt, coeff, order, max_res, C.byref(r), C.byref(n)))
        return r.value, n.value

    def diffusion_solve(self, dt, diffusion_coeff, order, max_res):
        return self._diffusion(0, dt, diffusion_coeff, order, max_res)

    def diffusion_solve_vcoeff(self, dt, order, max_res):
        return self._diffusion(1, dt, 1.0, order, max_res)

    def diffusion_solve_acoeff(self, dt, order, max_res):
        return self._diffusion(2, dt, 1.0, order, max_res)

    def diffusion_set_rhs(self, f1, f2):
        self._gck(self.G.omg_diffusion_set_rhs(self.gpu, f1, f2))

    def mg_poisson_free_3d(self, new_rhs, max_fft_frac, fmgcycle=True, want_res=True):
        """src/m_free_space.f90:36-214; returns max_res (None if not requested)"""
        r = C.c_double(0.0)
        self._ck(self.H.omgh_poisson_free_3d(self.h, int(bool(new_rhs)), max_fft_frac, int(bool(fmgcycle)), int(bool(want_res)), C.byref(r)))
        return r.value if want_res else None

    @property
    def free_space_fft_lvl(self):
        return self.H.omgh_free_space_fft_lvl(self.h)

    # coupling_amrvac/mod_multigrid_coupling.t:133-267: application arrays <-> device tree, no host cc arrays in between
    def copy_size(self, leaves_only=True, with_ghosts=False, lvl_lo=-20, lvl_hi=20):
        return int(self.G.omg_copy_size(self.gpu, lvl_lo, lvl_hi, int(leaves_only), int(with_ghosts)))

    def mg_copy_to_tree(self, iv, buf, leaves_only=True, with_ghosts=False, lvl_lo=-20, lvl_hi=20):
        buf = np.ascontiguousarray(buf, dtype=np.float64)
        self._gck(self.G.omg_copy_to_tree(self.gpu, iv, lvl_lo, lvl_hi, int(leaves_only), int(with_ghosts),
                                          buf.ctypes.data_as(C.c_void_p), buf.size))

    def mg_copy_from_tree(self, iv, leaves_only=True, with_ghosts=False, lvl_lo=-20, lvl_hi=20):
        out = np.empty(self.copy_size(leaves_only, with_ghosts, lvl_lo, lvl_hi))
        self._gck(self.G.omg_copy_from_tree(self.gpu, iv, lvl_lo, lvl_hi, int(leaves_only), int(with_ghosts),
                                            out.ctypes.data_as(C.c_void_p), out.size))
        return out

    def mg_gpu_push(self, iv):
        self._ck(self.H.omgh_gpu_push(self.h, iv))

    def mg_gpu_pull(self, iv, with_ghosts=True):
        self._ck(self.H.omgh_gpu_pull(self.h, iv, int(with_ghosts)))

    def mg_gpu_push_params(self):
        self._ck(self.H.omgh_gpu_push_params(self.h))

    def mg_gpu_push_bc(self):
        self._ck(self.H.omgh_gpu_push_bc(self.h))

    # -- C ABI conveniences ------------------------------------------------------------------------------
    def _gck(self, rc):
        if rc != 0:
            raise MgError(self.G.omg_last_error(self.gpu).decode())

    def stat(self, key):
        return int(self.G.omg_stat(self.gpu, key.encode()))

    def set_option(self, key, value):
        self._gck(self.G.omg_set_option(self.gpu, key.encode(), int(value)))

    def fas_vcycles(self, n, want_res=True):
        """n back-to-back device-resident V-cycles (omg_fas_vcycles); returns (residuals, elapsed_ms)."""
        res = (C.c_double * max(n, 1))()
        self._gck(self.G.omg_fas_vcycles(self.gpu, n, res if want_res else None))
        return (list(res[:n]) if want_res else None), float(self.G.omg_last_ms(self.gpu, b"cycles"))

    def trace(self, cap=100000):
        """steps recorded since set_option("trace", 1): list of (name, lvl, a, b, ms)"""
        rec = (C.c_int * (4 * cap))()
        ms = (C.c_double * cap)()
        n = C.c_int(0)
        self._gck(self.G.omg_trace(self.gpu, cap, rec, ms, C.byref(n)))
        return [(self.G.omg_step_name(rec[4 * i]).decode(), rec[4 * i + 1], rec[4 * i + 2], rec[4 * i + 3], ms[i])
                for i in range(min(n.value, cap))]

    def sync(self):
        self._gck(self.G.omg_sync(self.gpu))

    def mg_timers(self):
        """{name: seconds} of the reference's seven timers (mg_timers_show, src/m_data_structures.f90:562-581)."""
        n = C.c_int(0)
        self._gck(self.G.omg_timers(self.gpu, C.byref(n), None, None))
        names, secs = (C.c_char_p * n.value)(), (C.c_double * n.value)()
        self._gck(self.G.omg_timers(self.gpu, C.byref(n), names, secs))
        return {names[i].decode(): secs[i] for i in range(n.value)}

    # -- building blocks of the cycle (private in the reference; exposed for per-kernel parity tests) -------
    def _step(self, what, lvl, arg=0):
        """device-resident single step; in drop-in mode the four variables are pushed before and pulled after"""
        resident = bool(self.get_int("gpu_resident"))
        self._ck(self.H.omgh_gpu_push_params(self.h))
        if not resident:
            for iv in range(1, mg_num_vars + self.n_extra_vars + 1):
                self.mg_gpu_push(iv)
        out = C.c_double(0.0)
        self._gck(self.G.omg_debug_step(self.gpu, what.encode(), lvl, arg, C.byref(out)))
        if not resident:
            for iv in (1, 2, 3, 4):
                self.mg_gpu_pull(iv, True)
        return out.value

    def smooth_boxes(self, lvl, n_cycle):
        self._step("smooth_boxes", lvl, n_cycle)

    def update_coarse(self, lvl):
        self._step("update_coarse", lvl)

    def correct_children(self, lvl):
        self._step("correct_children", lvl)

    def max_residual_lvl(self, lvl):
        return self._step("max_residual_lvl", lvl)


class MgGroup:
    """n emulated ranks inside one process, for the parity tests on a single GPU (omg_comm_init_local).

    Every method call is forwarded to the Mg of every rank, one host thread per rank (the cycle's exchange
    steps meet at a host barrier); results must agree on all ranks.  `cc(id)` / `ids(lvl)` give the global
    view the oracle has, so that the same comparison helpers work."""

    def __init__(self, n_cpu, operator_type=mg_laplacian, smoother_type=mg_smoother_gsrb, n_extra_vars=0, device=0, **kw):
        self.n = n_cpu
        self.ranks = [Mg(operator_type, smoother_type, n_extra_vars=n_extra_vars, n_cpu=n_cpu, my_rank=r, device=device, **kw)
                      for r in range(n_cpu)]
        self._owner = None

    def _all(self, fn):
        import threading
        out, err = [None] * self.n, [None] * self.n

        def run(r):
            try:
                out[r] = fn(self.ranks[r])
            except BaseException as e:      # noqa: BLE001 - re-raised below
                err[r] = e
        ts = [threading.Thread(target=run, args=(r,)) for r in range(self.n)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        for e in err:
            if e is not None:
                raise e
        return out

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        attr = getattr(type(self.ranks[0]), name, None)
        if isinstance(attr, property):
            return getattr(self.ranks[0], name)

        def call(*a, **k):
            res = self._all(lambda m: getattr(m, name)(*a, **k))
            if isinstance(res[0], (int, float, str, list, tuple, type(None))):
                assert all(r == res[0] for r in res), "ranks disagree on %s: %r" % (name, res)
            return res[0]
        return call

    @property
    def highest_lvl(self):
        return self.ranks[0].highest_lvl

    @highest_lvl.setter
    def highest_lvl(self, v):
        for m in self.ranks:
            m.highest_lvl = v

    def mg_allocate_storage(self):
        for m in self.ranks:
            m.mg_allocate_storage()
        hs = (C.c_void_p * self.n)(*[m.gpu for m in self.ranks])
        if self.ranks[0].G.omg_comm_init_local(hs, self.n) != 0:
            raise MgError(self.ranks[0].G.omg_last_error(self.ranks[0].gpu).decode())
        self._owner = None

    def mg_gpu_update_tree(self):
        self._all(lambda m: m.mg_gpu_update_tree())
        hs = (C.c_void_p * self.n)(*[m.gpu for m in self.ranks])
        if self.ranks[0].G.omg_comm_init_local(hs, self.n) != 0:      # emulated ranks: the new plane pools are shared again
            raise MgError(self.ranks[0].G.omg_last_error(self.ranks[0].gpu).decode())
        self._owner = None

    def owner(self, id_):
        if self._owner is None:
            self._owner = self.ranks[0].tree_export()[:, 0]
        return self.ranks[int(self._owner[id_ - 1])]

    def cc(self, id_, nc=None):
        return self.owner(id_).cc(id_, nc)

    def my_ids(self, lvl):
        return self.ranks[0].ids(lvl)

    def ids(self, lvl):
        return self.ranks[0].ids(lvl)

    def leaves(self, lvl):
        return self.ranks[0].leaves(lvl)

    def stat(self, key):
        return sum(m.stat(key) for m in self.ranks)
