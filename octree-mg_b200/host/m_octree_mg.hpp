// m_octree_mg.hpp -- C++ host mirror of the reference's Fortran module surface (`use m_octree_mg`).
//
// The reference host side is Fortran 2008 (src/m_octree_mg.f90 re-exporting m_data_structures,
// m_build_tree, m_load_balance, m_allocate_storage, m_multigrid, m_ghost_cells, m_restrict,
// m_prolong).  This image has no Fortran compiler, so the host side above the C ABI is written in
// C++ with the same names, argument meaning and error behaviour; octree-mg_b200/fortran/ holds the
// ISO_C_BINDING shim a Fortran host links instead.  The data model (mg_t, mg_box_t, mg_lvl_t,
// mg_bc_t; src/m_data_structures.f90:181-342), tree construction and load balancing stay on the
// host exactly as in the reference; everything inside the cycle runs on the GPU through
// include/omg_gpu.h.  There is no CPU fallback.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/omg_gpu.h"

namespace omg {

// src/m_data_structures.f90:8-87
constexpr int mg_laplacian = 1, mg_vlaplacian = 2, mg_helmholtz = 3, mg_vhelmholtz = 4, mg_ahelmholtz = 5;
constexpr int mg_cartesian = 1, mg_cylindrical = 2, mg_spherical = 3;
constexpr int mg_smoother_gs = 1, mg_smoother_gsrb = 2, mg_smoother_jacobi = 3;
constexpr int mg_ndim = 3, mg_num_vars = 4, mg_max_num_vars = 10;
constexpr int mg_iphi = 1, mg_irhs = 2, mg_iold = 3, mg_ires = 4, mg_iveps = 5, mg_iveps1 = 5, mg_iveps2 = 6, mg_iveps3 = 7;
constexpr int mg_lvl_lo = -20, mg_lvl_hi = 20;
constexpr int mg_bc_dirichlet = -10, mg_bc_neumann = -11, mg_bc_continuous = -12;
constexpr int mg_no_box = 0, mg_physical_boundary = -1;
constexpr int mg_num_children = 8, mg_num_neighbors = 6;

// the reference `error stop`s; the mirror throws
struct mg_error : std::runtime_error { using std::runtime_error::runtime_error; };

struct mg_box_t {
  int rank = 0, id = 0, lvl = 0, ix[3] = {0, 0, 0}, parent = 0;
  int children[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int neighbors[6] = {0, 0, 0, 0, 0, 0};
  double r_min[3] = {0, 0, 0}, dr[3] = {0, 0, 0};
  std::vector<double> cc;   // cc(0:nc+1, 0:nc+1, 0:nc+1, nvar), i fastest; allocated for owned boxes only
};

struct mg_lvl_t { std::vector<int> leaves, parents, ref_bnds, ids, my_leaves, my_parents, my_ref_bnds, my_ids; };

// mg_subr_bc (src/m_data_structures.f90:344-360): bc is nc*nc, first tangential index fastest
typedef void (*mg_subr_bc)(const mg_box_t& box, int nc, int iv, int nb, int* bc_type, double* bc, void* user);

struct mg_bc_t {
  int bc_type = mg_bc_dirichlet;
  double bc_value = 0.0;
  mg_subr_bc boundary_cond = nullptr;
  void* user = nullptr;
};

struct mg_t {
  bool tree_created = false, is_allocated = false;
  int n_extra_vars = 0;
  int n_cpu = -1, my_rank = -1;
  int box_size = -1, highest_lvl = -1, lowest_lvl = -1, first_normal_lvl = -1, n_boxes = 0;
  int box_size_lvl[OMG_NLVL];
  int domain_size_lvl[OMG_NLVL][3];
  double dr[OMG_NLVL][3];
  double r_min[3] = {0, 0, 0};
  mg_lvl_t lvls[OMG_NLVL];
  std::vector<mg_box_t> boxes;     // boxes[id], ids are 1-based (boxes[0] unused)
  bool phi_bc_data_stored = false;
  bool periodic[3] = {false, false, false};
  mg_bc_t bc[mg_num_neighbors][mg_max_num_vars];
  int operator_type = mg_laplacian, geometry_type = mg_cartesian;
  bool subtract_mean = false;
  int smoother_type = mg_smoother_gs;
  int n_smoother_substeps = 1, n_cycle_down = 2, n_cycle_up = 2, max_coarse_cycles = 1000;
  int coarsest_grid[3] = {2, 2, 2};
  double residual_coarse_abs = 1e-8, residual_coarse_rel = 1e-8;
  bool methods_set = false;
  // module globals helmholtz_lambda / vhelmholtz_lambda / ahelmholtz_lambda (one per mg_t here)
  double lambda = 0.0;

  // ---- device side ------------------------------------------------------------------------
  omg_handle* gpu = nullptr;
  int device = 0;
  // false (default): strict drop-in, every mg_fas_* pushes phi/rhs(/eps) and pulls phi/rhs/old/res.
  // true: device-resident; the host calls mg_gpu_push / mg_gpu_pull around its own cc accesses.
  bool gpu_resident = false;
  bool bc_dirty = true, eps_dirty = true;
  int coarse_collapse_boxes = 0;   // > 0: coarse levels with fewer boxes per rank than this live on one GPU (see m_load_balance part)
  bool host_cc = true;   // false: device-resident runs that never touch host cc skip the host allocation
  // type(mg_free_bc_t) of src/m_free_space.f90:9-22: what the host keeps of it (grids and planes live on the device)
  bool free_bc_initialized = false;
  int free_bc_fft_lvl = -1;

  mg_lvl_t& L(int lvl) { return lvls[lvl - mg_lvl_lo]; }
  const mg_lvl_t& L(int lvl) const { return lvls[lvl - mg_lvl_lo]; }
  int nvar() const { return mg_num_vars + n_extra_vars; }
  mg_t();
  ~mg_t();
  mg_t(const mg_t&) = delete;
  mg_t& operator=(const mg_t&) = delete;
};

// ---- m_data_structures helpers (:430-539) --------------------------------------------------
bool mg_has_children(const mg_box_t& box);
int mg_ix_to_ichild(const int ix[3]);
void mg_get_child_offset(const mg_t& mg, int id, int off[3]);
int64_t mg_number_of_unknowns(const mg_t& mg);
void mg_get_face_coords(const mg_box_t& box, int nb, int nc, double* x /* [3][nc*nc] */);

// ---- m_communication: mg_comm_init (:12-35); ranks come from the launcher, NCCL from the C ABI ----
void mg_comm_init(mg_t& mg, int my_rank = 0, int n_cpu = 1);

// ---- m_build_tree ---------------------------------------------------------------------------
void mg_build_rectangle(mg_t& mg, const int domain_size[3], int box_size, const double dx[3], const double r_min[3],
                        const bool periodic[3], int n_finer);
void mg_add_children(mg_t& mg, int id);
void mg_set_leaves_parents(std::vector<mg_box_t>& boxes, mg_lvl_t& level);
void mg_set_next_level_ids(mg_t& mg, int lvl);
void mg_set_neighbors_lvl(mg_t& mg, int lvl);
void mg_set_refinement_boundaries(std::vector<mg_box_t>& boxes, mg_lvl_t& level);

// ---- m_load_balance ----------------------------------------------------------------------------
void mg_load_balance_simple(mg_t& mg);
void mg_load_balance(mg_t& mg);
void mg_load_balance_parents(mg_t& mg);

// ---- m_allocate_storage (host cc arrays + device mirrors) -----------------------------------------
void mg_allocate_storage(mg_t& mg);
void mg_deallocate_storage(mg_t& mg);
// after the host changed the tree of an allocated mg_t (refinement, re-partition): keep the device data of surviving boxes
void mg_gpu_update_tree(mg_t& mg);

// ---- m_multigrid / operator modules ------------------------------------------------------------
void mg_set_methods(mg_t& mg);
void helmholtz_set_lambda(mg_t& mg, double lambda);    // also vhelmholtz_ / ahelmholtz_set_lambda
void mg_fas_fmg(mg_t& mg, bool have_guess, double* max_res = nullptr);
void mg_fas_vcycle(mg_t& mg, const int* highest_lvl = nullptr, double* max_res = nullptr, bool standalone = true);
void mg_apply_op(mg_t& mg, int i_out);

// ---- m_diffusion (src/m_diffusion.f90:19-157): implicit diffusion steps; phi(t) in mg_iphi -> phi(t + dt) ----------
// Return the number of V-cycles added after the FMG cycle; *res (optional) receives the last max residual.
int diffusion_solve(mg_t& mg, double dt, double diffusion_coeff, int order, double max_res, double* res = nullptr);
int diffusion_solve_vcoeff(mg_t& mg, double dt, int order, double max_res, double* res = nullptr);   // eps in mg_iveps
int diffusion_solve_acoeff(mg_t& mg, double dt, int order, double max_res, double* res = nullptr);   // eps in mg_iveps1..3

// ---- m_free_space (src/m_free_space.f90:36-214): free-space boundary conditions for 3-D Poisson problems.  Each call
//      performs an additional FMG cycle (or V-cycle); max_fft_frac bounds the size of the level the FFT solve runs on ----
void mg_poisson_free_3d(mg_t& mg, bool new_rhs, double max_fft_frac, bool fmgcycle, double* max_res = nullptr);
int mg_highest_uniform_lvl(const mg_t& mg);

// ---- m_ghost_cells / m_restrict / m_prolong level-wide entry points ---------------------------------
void mg_fill_ghost_cells_lvl(mg_t& mg, int lvl, int iv);
void mg_fill_ghost_cells(mg_t& mg, int iv);
void mg_phi_bc_store(mg_t& mg);
void mg_restrict_lvl(mg_t& mg, int iv, int lvl);
void mg_restrict(mg_t& mg, int iv);
void mg_prolong(mg_t& mg, int lvl, int iv, int iv_to, bool add);

// ---- application arrays <-> device tree (mg_copy_to_tree / mg_copy_from_tree[_gc],
//      coupling_amrvac/mod_multigrid_coupling.t:133-267): `buf` holds the owned leaves (leaves_only) or all owned boxes of
//      every level in my_ids order, nc^3 interior cells per box or (nc+2)^3 with one ghost layer, i fastest ------------
int64_t mg_copy_size(const mg_t& mg, bool leaves_only = true, bool with_ghosts = false);
void mg_copy_to_tree(mg_t& mg, int iv, const double* buf, int64_t n_doubles, bool leaves_only = true, bool with_ghosts = false);
void mg_copy_from_tree(mg_t& mg, int iv, double* buf, int64_t n_doubles, bool leaves_only = true, bool with_ghosts = false);

// ---- host <-> device coherence ------------------------------------------------------
void mg_gpu_push(mg_t& mg, int iv);                       // all levels, owned boxes
void mg_gpu_pull(mg_t& mg, int iv, bool with_ghosts = true);
void mg_gpu_push_params(mg_t& mg);                        // mg_t scalars, bc constants (+ planes when dirty)
void mg_gpu_push_bc(mg_t& mg);                            // evaluate boundary_cond callbacks -> device planes

}  // namespace omg
