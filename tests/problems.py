"""Workload definitions shared by the oracle (CPU) and the product (GPU) -- mirrors of the reference's
test programs.  Every builder takes the class to instantiate (oracle.OracleMg or octree_mg_b200.Mg;
both expose the reference's routine names) so that the two sides get bit-identical inputs.

  uniform_problem      tests/test_uniform_grid.f90: manufactured solution prod sin(2 pi 5 x_d), rhs = the
                       discrete operator applied to it (ghost cells from the analytic field), Dirichlet
                       planes from a boundary_cond callback evaluating the solution at face centres.
  performance_problem  tests/test_performance.f90: rhs = 1, phi = 0, homogeneous Dirichlet constants.
  refinement_problem   tests/test_refinement.f90: centred refinement, solution restricted + ghost-filled
                       through the library itself, rhs = discrete operator, Dirichlet 0.
"""
import numpy as np

mg_laplacian, mg_vlaplacian, mg_helmholtz, mg_vhelmholtz, mg_ahelmholtz = 1, 2, 3, 4, 5
mg_smoother_gs, mg_smoother_gsrb = 1, 2
mg_iphi, mg_irhs, mg_iold, mg_ires, mg_iveps = 1, 2, 3, 4, 5
mg_num_vars = 4
mg_bc_dirichlet, mg_bc_neumann, mg_bc_continuous = -10, -11, -12

N_MODES = 5


def solution(x, y, z, n_modes=N_MODES):
    """tests/test_uniform_grid.f90:127-130"""
    return np.sin(2 * np.pi * n_modes * x) * np.sin(2 * np.pi * n_modes * y) * np.sin(2 * np.pi * n_modes * z)


def n_eps(operator_type):
    return {mg_laplacian: 0, mg_helmholtz: 0, mg_vlaplacian: 1, mg_vhelmholtz: 1, mg_ahelmholtz: 3}[operator_type]


def cell_coords(mg, id_, lvl):
    """Cell-centre coordinates incl. ghost cells, arrays [k, j, i] (tests/test_uniform_grid.f90:143-151)."""
    nc = mg.box_size_lvl(lvl)
    dr = mg.dr(lvl)
    r_min = mg.box_info(id_)["r_min"]
    idx = np.arange(0, nc + 2) - 0.5
    x = r_min[0] + idx * dr[0]
    y = r_min[1] + idx * dr[1]
    z = r_min[2] + idx * dr[2]
    Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    return X, Y, Z


def eps_field(X, Y, Z, d=0):
    """eps = max(1, 1 + x) as in tests/test_uniform_grid.f90:149; anisotropic variants vary per direction."""
    if d == 0:
        return np.maximum(1.0, 1.0 + X)
    if d == 1:
        return 1.0 + 0.5 * Y * Y
    return 2.0 - 0.5 * Z


def numpy_operator(phi, eps, operator_type, idr2, lam):
    """Discrete operator on one ghosted box [k, j, i] -> interior values.  Only used to manufacture the
    right-hand side (an input shared by both sides), so its rounding need not match anything."""
    c = phi[1:-1, 1:-1, 1:-1]
    nbr = [(phi[1:-1, 1:-1, :-2], phi[1:-1, 1:-1, 2:]), (phi[1:-1, :-2, 1:-1], phi[1:-1, 2:, 1:-1]),
           (phi[:-2, 1:-1, 1:-1], phi[2:, 1:-1, 1:-1])]
    if operator_type in (mg_laplacian, mg_helmholtz):
        out = sum(idr2[d] * (nbr[d][0] + nbr[d][1] - 2 * c) for d in range(3))
    else:
        out = np.zeros_like(c)
        for d in range(3):
            e = eps[d if operator_type == mg_ahelmholtz else 0]
            a0 = e[1:-1, 1:-1, 1:-1]
            lo = [e[1:-1, 1:-1, :-2], e[1:-1, :-2, 1:-1], e[:-2, 1:-1, 1:-1]][d]
            hi = [e[1:-1, 1:-1, 2:], e[1:-1, 2:, 1:-1], e[2:, 1:-1, 1:-1]][d]
            out = out + 2 * idr2[d] * a0 * lo / (a0 + lo) * (nbr[d][0] - c)
            out = out + 2 * idr2[d] * a0 * hi / (a0 + hi) * (nbr[d][1] - c)
    if operator_type in (mg_helmholtz, mg_vhelmholtz, mg_ahelmholtz):
        out = out - lam * c
    return out


def face_plane(fn, r_min, dr, nc, nb):
    """Evaluate fn(x, y, z) at the face centres of face nb -> plane [t2, t1] (mg_get_face_coords,
    src/m_data_structures.f90:495-539)."""
    d = (nb - 1) // 2
    low = (nb % 2) == 1
    t = np.arange(nc) + 0.5
    tdims = [q for q in range(3) if q != d]
    B, A = np.meshgrid(t, t, indexing="ij")      # plane[b, a]: a = first tangential index (fastest)
    c = [None] * 3
    c[tdims[0]] = r_min[tdims[0]] + A * dr[tdims[0]]
    c[tdims[1]] = r_min[tdims[1]] + B * dr[tdims[1]]
    c[d] = np.full_like(A, r_min[d] + (0.0 if low else dr[d] * nc))
    return fn(c[0], c[1], c[2])


def all_ids(mg):
    return [(lvl, id_) for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1) for id_ in mg.ids(lvl)]


def uniform_problem(cls, n, box, operator_type=mg_laplacian, smoother=mg_smoother_gsrb, bc="callback", lam=0.0,
                    periodic=(False, False, False), load_balance="default", setup=None, after_alloc=None, **kw):
    """test_uniform_grid semantics.  bc: 'callback' (Dirichlet planes of the exact solution),
    'dirichlet0', 'neumann0', 'mixed' (Neumann-0 everywhere, Dirichlet on -x), 'continuous_x'."""
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    ne = n_eps(operator_type)
    mg = cls(operator_type, smoother, n_extra_vars=ne + 1, **kw)
    i_sol = mg_num_vars + ne + 1
    if lam:
        mg.set_lambda(lam)
    for nb in range(1, 7):
        if bc == "callback":
            mg.set_bc_callback(nb, mg_iphi, lambda r_min, dr, lvl, nc, iv, nb_: (
                mg_bc_dirichlet, face_plane(solution, r_min, dr, nc, nb_)))
        elif bc == "dirichlet0":
            mg.set_bc(nb, mg_iphi, mg_bc_dirichlet, 0.0)
        elif bc == "neumann0":
            mg.set_bc(nb, mg_iphi, mg_bc_neumann, 0.0)
        elif bc == "mixed":
            mg.set_bc(nb, mg_iphi, mg_bc_dirichlet if nb == 1 else mg_bc_neumann, 0.25 if nb == 1 else 0.5)
        elif bc == "continuous_x":
            mg.set_bc(nb, mg_iphi, mg_bc_continuous if nb <= 2 else mg_bc_dirichlet, 0.0 if nb <= 2 else 1.0)
        else:
            raise ValueError(bc)
    mg.mg_set_methods()
    if setup is not None:          # e.g. host-mirror extensions that must be set before the partition is made
        setup(mg)
    dr = [1.0 / n3[0]] * 3
    mg.mg_build_rectangle(n3, box, dr, [0.0] * 3, list(periodic))
    if load_balance == "default":
        mg.mg_load_balance()
    elif load_balance == "simple":
        mg.mg_load_balance_simple()
    else:
        mg.mg_load_balance()
        mg.mg_load_balance_parents()
    mg.mg_allocate_storage()
    if after_alloc is not None:    # e.g. the communicator of a multi-process run (omg_comm_init follows omg_set_tree)
        after_alloc(mg)
    owned = hasattr(mg, "my_ids")
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        nc = mg.box_size_lvl(lvl)
        drl = mg.dr(lvl)
        idr2 = 1.0 / drl ** 2
        for id_ in (mg.my_ids(lvl) if owned else mg.ids(lvl)):
            cc = mg.cc(id_, nc)
            X, Y, Z = cell_coords(mg, id_, lvl)
            sol = solution(X, Y, Z)
            eps = [eps_field(X, Y, Z, d) for d in range(max(ne, 1))]
            for d in range(ne):
                cc[mg_iveps - 1 + d] = eps[d]
            cc[i_sol - 1] = sol
            cc[mg_irhs - 1, 1:-1, 1:-1, 1:-1] = numpy_operator(sol, eps, operator_type, idr2, lam)
            cc[mg_iphi - 1] = 0.0
    return mg, i_sol


def performance_problem(cls, n, box, smoother=mg_smoother_gsrb, **kw):
    """tests/test_performance.f90:51-66,102-115"""
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    mg = cls(mg_laplacian, smoother, **kw)
    for nb in range(1, 7):
        mg.set_bc(nb, mg_iphi, mg_bc_dirichlet, 0.0)
    mg.mg_set_methods()
    mg.mg_build_rectangle(n3, box, [1.0 / n3[0]] * 3, [0.0] * 3, [False] * 3)
    mg.mg_load_balance()
    mg.mg_load_balance_parents()
    mg.mg_allocate_storage()
    owned = hasattr(mg, "my_ids")
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        nc = mg.box_size_lvl(lvl)
        for id_ in (mg.my_ids(lvl) if owned else mg.ids(lvl)):
            mg.cc(id_, nc)[mg_irhs - 1, 1:-1, 1:-1, 1:-1] = 1.0
    return mg


def build_amr_tree(mg, n_amr_levels, lvl1_size, box_size, dr, periodic=(False, False, False), refine=None):
    """tests/test_refinement.f90:191-247.  refine(lvl, box_center, domain_len) -> bool overrides the
    default centred-region rule."""
    lvl1_size = list(lvl1_size)
    n_finer = n_amr_levels * int(np.prod([s // box_size for s in lvl1_size])) * 8 + 1000
    domain_len = np.array(lvl1_size) * np.array(dr)
    mg.mg_build_rectangle(lvl1_size, box_size, list(dr), [0.0] * 3, list(periodic), n_finer)
    for lvl in range(1, n_amr_levels):
        for id_ in mg.ids(lvl):
            info = mg.box_info(id_)
            r0 = 0.5 * domain_len - domain_len * 0.5 ** (lvl + 1)
            r1 = 0.5 * domain_len + domain_len * 0.5 ** (lvl + 1)
            center = info["r_min"] + 0.5 * box_size * mg.dr(lvl)
            if refine is not None:
                do = refine(lvl, center, domain_len)
            else:
                do = bool(np.all((center >= r0) & (center <= r1)))
            if do:
                mg.mg_add_children(id_)
        mg.mg_set_leaves_parents(lvl)
        mg.mg_set_next_level_ids(lvl)
        mg.mg_set_neighbors_lvl(lvl + 1)
    mg.mg_set_leaves_parents(n_amr_levels)
    mg.highest_lvl = n_amr_levels
    for lvl in range(1, mg.highest_lvl + 1):
        mg.mg_set_refinement_boundaries(lvl)
    mg.mg_load_balance()
    mg.mg_allocate_storage()


def refinement_problem(cls, n_levels, box, n, smoother=mg_smoother_gsrb, refine=None, **kw):
    """tests/test_refinement.f90 (Laplacian, Dirichlet 0).  The solution's ghost cells come from the
    library's own restriction + ghost-cell fill, as in set_solution (:124-145); the product runs those
    on the device through phi (whose ghost cells it derives), the oracle through i_sol like the reference."""
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    mg = cls(mg_laplacian, smoother, n_extra_vars=1, **kw)
    i_sol = mg_num_vars + 1
    for nb in range(1, 7):
        mg.set_bc(nb, mg_iphi, mg_bc_dirichlet, 0.0)
        mg.set_bc(nb, i_sol, mg_bc_dirichlet, 0.0)
    mg.mg_set_methods()
    build_amr_tree(mg, n_levels, n3, box, [1.0 / n3[0]] * 3, refine=refine)
    owned = hasattr(mg, "my_ids")
    boxes = [(lvl, id_) for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1)
             for id_ in (mg.my_ids(lvl) if owned else mg.ids(lvl))]
    for lvl, id_ in boxes:
        X, Y, Z = cell_coords(mg, id_, lvl)
        mg.cc(id_)[i_sol - 1] = solution(X, Y, Z)
    mg.mg_restrict(i_sol)
    mg.mg_fill_ghost_cells(i_sol)
    for lvl, id_ in boxes:
        cc = mg.cc(id_)
        cc[mg_iphi - 1] = cc[i_sol - 1]
    if hasattr(mg, "box_op"):      # oracle: per-box operator exactly like compute_rhs_and_reset (:147-163)
        for lvl, id_ in boxes:
            mg.box_op(id_, mg_irhs)
    else:                          # product: ghost cells of phi are derived on the device
        mg.mg_fill_ghost_cells(mg_iphi)
        mg.mg_apply_op(mg_irhs)
    for lvl, id_ in boxes:
        mg.cc(id_)[mg_iphi - 1] = 0.0
    return mg, i_sol


DIFF_MODES = (1, 2, 3)


def diffusion_mode(X, Y, Z):
    """a discrete eigenfunction of the cell-centred 7-point Laplacian with homogeneous Dirichlet faces on [0,1]^3"""
    return np.sin(np.pi * DIFF_MODES[0] * X) * np.sin(np.pi * DIFF_MODES[1] * Y) * np.sin(np.pi * DIFF_MODES[2] * Z)


def diffusion_eigenvalue(n3):
    """sum_d (4/h_d^2) sin^2(pi m_d h_d / 2) >= 0: -Laplacian_h(mode) = this * mode"""
    return sum(4.0 * n ** 2 * np.sin(np.pi * m / (2.0 * n)) ** 2 for n, m in zip(n3, DIFF_MODES))


def diffusion_problem(cls, n, box, operator_type=mg_helmholtz, amr_levels=0, **kw):
    """input of the solvers of src/m_diffusion.f90: phi(t) on every box, homogeneous Dirichlet faces, the
    diffusion coefficient(s) of the variable-coefficient solvers in mg_iveps.. on all levels incl. ghost cells"""
    n3 = [n] * 3 if np.isscalar(n) else list(n)
    ne = n_eps(operator_type)
    mg = cls(operator_type, mg_smoother_gsrb, n_extra_vars=ne, **kw)
    for nb in range(1, 7):
        mg.set_bc(nb, mg_iphi, mg_bc_dirichlet, 0.0)
    mg.mg_set_methods()
    if amr_levels:
        build_amr_tree(mg, amr_levels, n3, box, [1.0 / n3[0]] * 3)
    else:
        mg.mg_build_rectangle(n3, box, [1.0 / n3[0]] * 3, [0.0] * 3, [False] * 3)
        mg.mg_load_balance()
        mg.mg_allocate_storage()
    owned = hasattr(mg, "my_ids")
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        nc = mg.box_size_lvl(lvl)
        for id_ in (mg.my_ids(lvl) if owned else mg.ids(lvl)):
            cc = mg.cc(id_, nc)
            X, Y, Z = cell_coords(mg, id_, lvl)
            cc[mg_iphi - 1] = diffusion_mode(X, Y, Z)
            for d in range(ne):
                cc[mg_iveps - 1 + d] = eps_field(X, Y, Z, d)
    # The order-2 solvers apply the operator to phi(t) first (m_diffusion.f90:36-38), which reads phi's ghost cells:
    # make them the ones a ghost-cell fill produces, as they are after any previous solve.  (The device never stores
    # phi ghost cells -- it derives them, DESIGN.md section 3 -- so for it this is the only possible state.)
    mg.mg_fill_ghost_cells(mg_iphi)
    return mg


def leaf_phi(mg, iv=mg_iphi):
    """{id: interior of variable iv} over all leaves (copy)."""
    out = {}
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        for id_ in mg.leaves(lvl):
            out[id_] = mg.cc(id_)[iv - 1, 1:-1, 1:-1, 1:-1].copy()
    return out


def max_error(mg, i_sol, leaves_only=False):
    """print_error of the test programs: max |phi - sol| on the finest level (uniform) or on leaves (AMR)."""
    err = 0.0
    lvls = range(1, mg.highest_lvl + 1) if leaves_only else [mg.highest_lvl]
    for lvl in lvls:
        for id_ in (mg.leaves(lvl) if leaves_only else mg.ids(lvl)):
            cc = mg.cc(id_)
            err = max(err, float(np.abs(cc[mg_iphi - 1, 1:-1, 1:-1, 1:-1] - cc[i_sol - 1, 1:-1, 1:-1, 1:-1]).max()))
    return err


def run_cycles(mg, n_its, fmg):
    """The iteration loop of the test programs (tests/test_uniform_grid.f90:98-107); returns max_res per cycle."""
    hist = []
    for it in range(1, n_its + 1):
        hist.append(mg.mg_fas_fmg(it > 1) if fmg else mg.mg_fas_vcycle())
    return hist
