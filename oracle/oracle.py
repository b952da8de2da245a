"""ctypes wrapper around oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

The oracle is the CPU restatement of the reference's multigrid hot path
(oracle/omg_oracle.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product
package (octree-mg_b200/) must never do so.

The class mirrors the reference's Fortran API names (mg_build_rectangle,
mg_load_balance, mg_fas_fmg, mg_fas_vcycle, ...; src/m_octree_mg.f90).  Box data
is exposed as numpy views of cc with shape (nvar, nc+2, nc+2, nc+2) indexed
[v, k, j, i] (the Fortran cc(i, j, k, v+1)).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# constants of src/m_data_structures.f90:8-87
mg_laplacian, mg_vlaplacian, mg_helmholtz, mg_vhelmholtz, mg_ahelmholtz = 1, 2, 3, 4, 5
mg_smoother_gs, mg_smoother_gsrb = 1, 2
mg_iphi, mg_irhs, mg_iold, mg_ires, mg_iveps = 1, 2, 3, 4, 5
mg_num_vars = 4
mg_bc_dirichlet, mg_bc_neumann, mg_bc_continuous = -10, -11, -12
mg_num_neighbors = 6

BC_CALLBACK = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int,
                          C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_void_p)


def build():
    """Compile liboracle.so (g++ -O2 -ffp-contract=off)."""
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_create.restype = C.c_void_p
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_box_cc.restype = C.POINTER(C.c_double)
        L.orc_box_cc.argtypes = [C.c_void_p, C.c_int]
        L.orc_timer.restype = C.c_double
        L.orc_timer.argtypes = [C.c_void_p, C.c_int]
        L.orc_set_real.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.orc_set_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.orc_get_int.argtypes = [C.c_void_p, C.c_char_p]
        L.orc_set_bc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double]
        L.orc_set_bc_callback.argtypes = [C.c_void_p, C.c_int, C.c_int, BC_CALLBACK, C.c_void_p]
        for name in ("orc_destroy", "orc_set_methods", "orc_allocate_storage", "orc_phi_bc_store",
                     "orc_update_lvl_info"):
            getattr(L, name).argtypes = [C.c_void_p]
        for name in ("orc_add_children", "orc_set_leaves_parents", "orc_set_next_level_ids", "orc_set_neighbors_lvl",
                     "orc_set_refinement_boundaries", "orc_load_balance", "orc_fill_ghost_cells", "orc_restrict",
                     "orc_apply_op", "orc_box_size_lvl", "orc_update_coarse", "orc_correct_children"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_int]
        for name in ("orc_fill_ghost_cells_lvl", "orc_restrict_lvl", "orc_box_op", "orc_smooth_boxes",
                     "orc_set_box_rank"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_prolong.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_build_rectangle.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_double),
                                          C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]
        L.orc_lvl_list.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.orc_box_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.orc_tree_export.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.orc_dr_lvl.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        L.orc_fas_vcycle.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_int]
        L.orc_fas_fmg.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.orc_max_residual_lvl.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        L.orc_diffusion_set_rhs.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.orc_diffusion_solve.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, C.c_double,
                                          C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.orc_set_omp_threads.argtypes = [C.c_int]
        L.orc_poisson_free_3d.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.orc_free_space_fft_lvl.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


class OracleError(RuntimeError):
    pass


def _arr(vals, ctype):
    return (ctype * len(vals))(*vals)


class OracleMg:
    """CPU oracle with the reference's mg_t surface."""

    def __init__(self, operator_type=mg_laplacian, smoother_type=mg_smoother_gsrb, n_extra_vars=0, n_cpu=1):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_create())
        self._keep = []
        self.set_int("operator_type", operator_type)
        self.set_int("smoother_type", smoother_type)
        self.set_int("n_extra_vars", n_extra_vars)
        self.set_int("n_cpu", n_cpu)

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise OracleError(self.L.orc_last_error(self.h).decode())

    # -- parameters ---------------------------------------------------------
    def set_int(self, key, v):
        self._ck(self.L.orc_set_int(self.h, key.encode(), int(v)))

    def get_int(self, key):
        return self.L.orc_get_int(self.h, key.encode())

    def set_real(self, key, v):
        self._ck(self.L.orc_set_real(self.h, key.encode(), float(v)))

    def set_lambda(self, lam):
        """helmholtz_set_lambda / vhelmholtz_set_lambda / ahelmholtz_set_lambda."""
        self.set_real("lambda", lam)

    def set_bc(self, nb, iv, bc_type, bc_value=0.0):
        self._ck(self.L.orc_set_bc(self.h, nb, iv, bc_type, bc_value))

    def set_bc_callback(self, nb, iv, fn):
        """fn(r_min[3], dr[3], lvl, nc, iv, nb) -> (bc_type, bc[nc, nc] indexed [t2, t1])."""
        def tramp(r_min, dr, lvl, nc, iv_, nb_, bc_type, bc, user):
            t, plane = fn(np.array(r_min[0:3]), np.array(dr[0:3]), lvl, nc, iv_, nb_)
            bc_type[0] = t
            out = np.ctypeslib.as_array(bc, shape=(nc, nc))
            out[...] = plane
        cb = BC_CALLBACK(tramp)
        self._keep.append(cb)
        self._ck(self.L.orc_set_bc_callback(self.h, nb, iv, cb, None))

    # -- tree / setup -------------------------------------------------------
    def mg_set_methods(self):
        self._ck(self.L.orc_set_methods(self.h))

    def mg_build_rectangle(self, domain_size, box_size, dx, r_min, periodic, n_finer=0):
        self._ck(self.L.orc_build_rectangle(self.h, _arr(list(domain_size), C.c_int), box_size,
                                            _arr(list(dx), C.c_double), _arr(list(r_min), C.c_double),
                                            _arr([int(bool(p)) for p in periodic], C.c_int), n_finer))

    def mg_add_children(self, id_):
        self._ck(self.L.orc_add_children(self.h, id_))

    def mg_set_leaves_parents(self, lvl):
        self._ck(self.L.orc_set_leaves_parents(self.h, lvl))

    def mg_set_next_level_ids(self, lvl):
        self._ck(self.L.orc_set_next_level_ids(self.h, lvl))

    def mg_set_neighbors_lvl(self, lvl):
        self._ck(self.L.orc_set_neighbors_lvl(self.h, lvl))

    def mg_set_refinement_boundaries(self, lvl):
        self._ck(self.L.orc_set_refinement_boundaries(self.h, lvl))

    def mg_load_balance_simple(self):
        self._ck(self.L.orc_load_balance(self.h, 0))

    def mg_load_balance(self):
        self._ck(self.L.orc_load_balance(self.h, 1))

    def mg_load_balance_parents(self):
        self._ck(self.L.orc_load_balance(self.h, 2))

    def mg_allocate_storage(self):
        self._ck(self.L.orc_allocate_storage(self.h))

    # -- introspection ------------------------------------------------------
    @property
    def lowest_lvl(self):
        return self.get_int("lowest_lvl")

    @property
    def highest_lvl(self):
        return self.get_int("highest_lvl")

    @highest_lvl.setter
    def highest_lvl(self, v):
        self.set_int("highest_lvl", v)

    @property
    def first_normal_lvl(self):
        return self.get_int("first_normal_lvl")

    @property
    def n_boxes(self):
        return self.get_int("n_boxes")

    @property
    def box_size(self):
        return self.get_int("box_size")

    @property
    def n_extra_vars(self):
        return self.get_int("n_extra_vars")

    def box_size_lvl(self, lvl):
        return self.L.orc_box_size_lvl(self.h, lvl)

    def dr(self, lvl):
        out = (C.c_double * 3)()
        self.L.orc_dr_lvl(self.h, lvl, out)
        return np.array(out[:])

    def _lst(self, lvl, which):
        n = self.L.orc_lvl_list(self.h, lvl, which, None)
        out = (C.c_int * max(n, 1))()
        self.L.orc_lvl_list(self.h, lvl, which, out)
        return list(out[:n])

    def ids(self, lvl):
        return self._lst(lvl, 0)

    def leaves(self, lvl):
        return self._lst(lvl, 1)

    def parents(self, lvl):
        return self._lst(lvl, 2)

    def ref_bnds(self, lvl):
        return self._lst(lvl, 3)

    def box_info(self, id_):
        out = (C.c_int * 20)()
        rmin = (C.c_double * 3)()
        self.L.orc_box_info(self.h, id_, out, rmin)
        o = list(out)
        return dict(rank=o[0], lvl=o[1], ix=o[2:5], parent=o[5], children=o[6:14], neighbors=o[14:20],
                    r_min=np.array(rmin[:]))

    def tree_export(self):
        """(n_boxes, 20) int32: rank, lvl, ix(3), parent, children(8), neighbors(6) for ids 1..n_boxes."""
        n = self.n_boxes
        out = np.zeros((n, 20), dtype=np.int32)
        self.L.orc_tree_export(self.h, out.ctypes.data_as(C.POINTER(C.c_int)))
        return out

    def set_box_rank(self, id_, rank):
        self.L.orc_set_box_rank(self.h, id_, rank)

    def update_lvl_info(self):
        self._ck(self.L.orc_update_lvl_info(self.h))

    def cc(self, id_, nc=None):
        """numpy view [v, k, j, i] of boxes(id)%cc (ghost cells included)."""
        if nc is None:
            nc = self.box_size_lvl(self.box_info(id_)["lvl"])
        nv = mg_num_vars + self.n_extra_vars
        p = self.L.orc_box_cc(self.h, id_)
        return np.ctypeslib.as_array(p, shape=(nv, nc + 2, nc + 2, nc + 2))

    # -- hot path -----------------------------------------------------------
    def mg_fill_ghost_cells_lvl(self, lvl, iv):
        self._ck(self.L.orc_fill_ghost_cells_lvl(self.h, lvl, iv))

    def mg_fill_ghost_cells(self, iv):
        self._ck(self.L.orc_fill_ghost_cells(self.h, iv))

    def mg_restrict_lvl(self, iv, lvl):
        self._ck(self.L.orc_restrict_lvl(self.h, iv, lvl))

    def mg_restrict(self, iv):
        self._ck(self.L.orc_restrict(self.h, iv))

    def mg_prolong(self, lvl, iv, iv_to, add):
        self._ck(self.L.orc_prolong(self.h, lvl, iv, iv_to, int(bool(add))))

    def mg_apply_op(self, i_out):
        self._ck(self.L.orc_apply_op(self.h, i_out))

    def box_op(self, id_, i_out):
        self._ck(self.L.orc_box_op(self.h, id_, i_out))

    def mg_phi_bc_store(self):
        self._ck(self.L.orc_phi_bc_store(self.h))

    def smooth_boxes(self, lvl, n_cycle):
        self._ck(self.L.orc_smooth_boxes(self.h, lvl, n_cycle))

    def update_coarse(self, lvl):
        self._ck(self.L.orc_update_coarse(self.h, lvl))

    def correct_children(self, lvl):
        self._ck(self.L.orc_correct_children(self.h, lvl))

    def max_residual_lvl(self, lvl):
        r = C.c_double(0.0)
        self._ck(self.L.orc_max_residual_lvl(self.h, lvl, C.byref(r)))
        return r.value

    def mg_fas_vcycle(self, highest_lvl=None, want_res=True, standalone=True):
        r = C.c_double(0.0)
        has = highest_lvl is not None
        self._ck(self.L.orc_fas_vcycle(self.h, int(has), highest_lvl if has else 0, int(want_res), C.byref(r),
                                       int(standalone)))
        return r.value if want_res else None

    def mg_fas_fmg(self, have_guess, want_res=True):
        r = C.c_double(0.0)
        self._ck(self.L.orc_fas_fmg(self.h, int(bool(have_guess)), int(want_res), C.byref(r)))
        return r.value if want_res else None

    # src/m_diffusion.f90: diffusion_solve (:19-58), diffusion_solve_vcoeff (:64-106), diffusion_solve_acoeff (:113-157)
    def _diffusion(self, op, dt, coeff, order, max_res):
        r, n = C.c_double(0.0), C.c_int(0)
        self._ck(self.L.orc_diffusion_solve(self.h, op, dt, coeff, order, max_res, C.byref(r), C.byref(n)))
        return r.value, n.value

    def diffusion_solve(self, dt, diffusion_coeff, order, max_res):
        return self._diffusion(3, dt, diffusion_coeff, order, max_res)

    def diffusion_solve_vcoeff(self, dt, order, max_res):
        return self._diffusion(4, dt, 1.0, order, max_res)

    def diffusion_solve_acoeff(self, dt, order, max_res):
        return self._diffusion(5, dt, 1.0, order, max_res)

    def diffusion_set_rhs(self, f1, f2):
        self._ck(self.L.orc_diffusion_set_rhs(self.h, f1, f2))

    def mg_poisson_free_3d(self, new_rhs, max_fft_frac, fmgcycle=True, want_res=True):
        """src/m_free_space.f90:36-214; returns max_res (None if not requested)"""
        r = C.c_double(0.0)
        self._ck(self.L.orc_poisson_free_3d(self.h, int(bool(new_rhs)), max_fft_frac, int(bool(fmgcycle)), int(bool(want_res)), C.byref(r)))
        return r.value if want_res else None

    @property
    def free_space_fft_lvl(self):
        return self.L.orc_free_space_fft_lvl(self.h)

    def timers(self):
        names = ["mg total V-cycle", "mg total FMG cycle", "mg smoother", "mg smoother g.c.", "mg coarse",
                 "mg correct", "mg update coarse"]
        return {n: self.L.orc_timer(self.h, i) for i, n in enumerate(names)}
