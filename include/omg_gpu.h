/* include/omg_gpu.h -- C ABI of libomg_b200.so
 *
 * Drop-in boundary for octree-mg's multigrid-cycle hot path on NVIDIA B200
 * (sm_100a).  A host (the Fortran shim in octree-mg_b200/fortran/, the C++ host
 * mirror in octree-mg_b200/host/, or any FFI) keeps the reference's data model
 * (mg_t, tree construction, load balancing, storage) and
 * calls these entry points instead of the bodies of src/m_multigrid.f90 and the
 * level-wide routines of src/m_ghost_cells.f90 / m_restrict.f90 / m_prolong.f90.
 *
 * Conventions
 *  - plain C, no torch / C++ types; every function returns 0 on success and a
 *    non-zero code on failure, omg_last_error() gives the message (the
 *    reference `error stop`s; the shim converts codes back to error stop).
 *  - box ids, variable indices (iv) and neighbour indices (nb) are 1-based
 *    exactly as in the reference (mg_iphi=1, mg_irhs=2, mg_iold=3, mg_ires=4,
 *    mg_iveps=5..; nb 1..6 = -x,+x,-y,+y,-z,+z).
 *  - a host box variable is the reference's contiguous cc(0:nc+1,0:nc+1,0:nc+1,iv)
 *    slice: (nc+2)^3 doubles, i fastest (src/m_data_structures.f90:206).
 *  - there is no CPU fallback: without a CUDA device every call fails loudly.
 */
#ifndef OMG_GPU_H
#define OMG_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OMG_LVL_LO (-20)   /* mg_lvl_lo, src/m_data_structures.f90:62 */
#define OMG_LVL_HI 20      /* mg_lvl_hi, :64 */
#define OMG_NLVL 41
#define OMG_MAX_VARS 10    /* mg_max_num_vars, :36 */

typedef struct omg_handle omg_handle;

/* Flattened view of type(mg_t) tree metadata (src/m_data_structures.f90:250-342,
 * per-box fields :194-207, level lists :181-191).  All arrays are read during the
 * call only.  Per-box arrays are indexed [id-1]. */
typedef struct omg_tree_desc {
  int32_t n_boxes;              /* mg%n_boxes */
  int32_t box_size;             /* mg%box_size */
  int32_t lowest_lvl, highest_lvl, first_normal_lvl;
  int32_t my_rank, n_cpu;       /* mg%my_rank, mg%n_cpu */
  int32_t n_extra_vars;         /* mg%n_extra_vars */
  int32_t periodic[3];          /* mg%periodic */
  const int32_t* box_size_lvl;  /* [OMG_NLVL], index lvl - OMG_LVL_LO */
  const double*  dr;            /* [OMG_NLVL][3] */
  const int32_t* box_rank;      /* [n_boxes]    boxes(id)%rank */
  const int32_t* box_lvl;       /* [n_boxes]    boxes(id)%lvl */
  const int32_t* box_ix;        /* [n_boxes][3] boxes(id)%ix */
  const int32_t* box_parent;    /* [n_boxes]    boxes(id)%parent */
  const int32_t* box_children;  /* [n_boxes][8] boxes(id)%children */
  const int32_t* box_neighbors; /* [n_boxes][6] boxes(id)%neighbors (>0 id, 0 refinement bnd, <0 physical) */
  const int32_t* lvl_ptr;       /* [OMG_NLVL+1] CSR offsets into lvl_ids, index lvl - OMG_LVL_LO */
  const int32_t* lvl_ids;       /* mg%lvls(lvl)%ids in the reference's order */
} omg_tree_desc;

/* Cycle parameters read by mg_set_methods / mg_fas_* (src/m_data_structures.f90:300-327,
 * lambda: src/m_helmholtz.f90:11, m_vhelmholtz.f90:12, m_ahelmholtz.f90:12). */
typedef struct omg_params {
  int32_t operator_type;        /* mg_laplacian=1 .. mg_ahelmholtz=5 */
  int32_t smoother_type;        /* mg_smoother_gs=1, mg_smoother_gsrb=2 */
  int32_t geometry_type;        /* mg_cartesian=1 only */
  int32_t n_cycle_down, n_cycle_up, max_coarse_cycles;
  int32_t subtract_mean;        /* mg%subtract_mean after mg_set_methods / mg_build_rectangle */
  double  residual_coarse_abs, residual_coarse_rel;
  double  lambda;               /* helmholtz_lambda / vhelmholtz_lambda / ahelmholtz_lambda */
} omg_params;

/* ---- lifetime ------------------------------------------------------------ */
int omg_create(int device, omg_handle** out);             /* one handle per mg_t, bound to one GPU */
int omg_destroy(omg_handle* h);                           /* mg_deallocate_storage, src/m_allocate_storage.f90:14-47 */
/* On-device regrid / re-partition.  The reference changes a tree with mg_add_children / mg_set_refinement_boundaries
 * (src/m_build_tree.f90:296-367) and mg_load_balance (src/m_load_balance.f90:71-136); its application coupling then
 * deallocates, rebuilds and copies every variable in again.  omg_update_tree takes the NEW description (same box size,
 * variables and ranks) and keeps the device data of every box that exists in both trees on the same level: boxes that
 * stay on this rank are copied device to device, boxes whose owner changed travel old owner -> new owner (NCCL send /
 * recv), new boxes are zero-filled as mg_allocate_storage does (src/m_allocate_storage.f90:72-76).  Collective on several
 * ranks.  Boundary planes, parameters and (for emulated ranks) the communicator's pool mappings are set again by the
 * caller as after omg_set_tree. */
int omg_update_tree(omg_handle* h, const omg_tree_desc* tree);
const char* omg_last_error(const omg_handle* h);          /* message of the last failing call */
const char* omg_version(void);

/* ---- setup (replaces what mg_allocate_storage sizes, src/m_allocate_storage.f90:51-99) */
int omg_set_tree(omg_handle* h, const omg_tree_desc* t);  /* device mirrors, all variables zero-filled (:72-76) */
int omg_set_params(omg_handle* h, const omg_params* p);   /* mg_set_methods, src/m_multigrid.f90:27-60 */

/* Physical boundary conditions of variable iv (mg%bc(nb, iv), src/m_data_structures.f90:232-242).
 * omg_set_bc_const: bc_type/bc_value constants.  omg_set_bc_plane: the plane a
 * boundary_cond callback (or mg_phi_bc_store, src/m_ghost_cells.f90:66-117) yields for one
 * boundary face of one box: nc*nc doubles, first tangential index fastest, plus its bc_type. */
int omg_set_bc_const(omg_handle* h, int iv, int nb, int bc_type, double bc_value);
int omg_set_bc_plane(omg_handle* h, int box_id, int nb, int bc_type, const double* plane);

/* ---- host <-> device coherence (hosts read/write mg%boxes(id)%cc directly) */
/* ptrs[i] -> cc(:,:,:,iv) of box ids[i] ((nc+2)^3 doubles).  push reads interior cells (and,
 * for iv >= 5, the six ghost faces, which hold user data such as eps); pull writes interior
 * cells and, when with_ghosts != 0, the six ghost faces as the reference would hold them. */
int omg_push_var(omg_handle* h, int iv, int n, const int32_t* ids, const double* const* ptrs);
int omg_pull_var(omg_handle* h, int iv, int n, const int32_t* ids, double* const* ptrs, int with_ghosts);
/* Same, for all owned boxes of levels lvl_lo..lvl_hi from/to one packed host buffer (boxes in
 * my_ids order, each (nc+2)^3 doubles); intended for pinned staging buffers. */
int omg_push_var_packed(omg_handle* h, int iv, int lvl_lo, int lvl_hi, const double* buf, int64_t n_doubles);
int omg_pull_var_packed(omg_handle* h, int iv, int lvl_lo, int lvl_hi, double* buf, int64_t n_doubles, int with_ghosts);
int64_t omg_packed_size(omg_handle* h, int lvl_lo, int lvl_hi);  /* doubles needed by the packed calls */
int omg_set_var_const(omg_handle* h, int iv, int lvl_lo, int lvl_hi, double value, int interior_only);
/* mg_copy_to_tree / mg_copy_from_tree / mg_copy_from_tree_gc of the reference's application coupling
 * (coupling_amrvac/mod_multigrid_coupling.t:133-171, 206-234, 238-267): application data <-> one variable of the tree,
 * straight between a host buffer and the device tree.  The buffer holds the owned boxes of levels lvl_lo..lvl_hi in
 * my_ids order (leaves_only = 1: the owned leaves, as the coupling copies grid leaves), per box nc^3 interior cells
 * (with_ghosts = 0) or (nc+2)^3 cells incl. one ghost layer (with_ghosts = 1), i fastest.  omg_copy_size = doubles needed. */
int omg_copy_to_tree(omg_handle* h, int iv, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts, const double* buf, int64_t n_doubles);
int omg_copy_from_tree(omg_handle* h, int iv, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts, double* buf, int64_t n_doubles);
int64_t omg_copy_size(omg_handle* h, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts);

/* ---- the hot path -------------------------------------------------------- */
/* mg_fas_vcycle(mg, highest_lvl, max_res, standalone), src/m_multigrid.f90:150-243.
 * Optional Fortran arguments become explicit flags. */
int omg_fas_vcycle(omg_handle* h, int has_highest_lvl, int highest_lvl, int standalone,
                   int want_res, double* max_res);
/* mg_fas_fmg(mg, have_guess, max_res), src/m_multigrid.f90:84-147 */
int omg_fas_fmg(omg_handle* h, int have_guess, int want_res, double* max_res);
/* mg_fill_ghost_cells_lvl, src/m_ghost_cells.f90:131-175 (refreshes device ghost state of iv) */
int omg_fill_ghost_cells_lvl(omg_handle* h, int lvl, int iv);
/* mg_restrict_lvl, src/m_restrict.f90:83-114 */
int omg_restrict_lvl(omg_handle* h, int iv, int lvl);
/* mg_prolong with the default mg_prolong_sparse, src/m_prolong.f90:51-85,159-240 */
int omg_prolong(omg_handle* h, int lvl, int iv, int iv_to, int add);
/* mg_apply_op with the built-in operator, src/m_multigrid.f90:439-456 */
int omg_apply_op(omg_handle* h, int i_out);
/* set_rhs of the implicit diffusion solvers, src/m_diffusion.f90:159-173: rhs <- f1 * phi + f2 * rhs on the owned
 * leaves of levels 1..highest.  diffusion_solve / _vcoeff / _acoeff (:19-157) are this call between omg_set_params
 * (lambda), omg_apply_op (order 2), omg_fas_fmg and omg_fas_vcycle: nothing but the residual leaves the device. */
int omg_diffusion_set_rhs(omg_handle* h, double f1, double f2);
/* Free-space front end, mg_poisson_free_3d (src/m_free_space.f90:36-214).  The reference solves the free-space problem
 * on a coarse fully refined level with the vendored BigDFT solver (createKernel :119, PSolver :160) and turns the result
 * into Dirichlet planes (ghost_cells_free_bc / interp_bc :217-270) and an initial guess.  Device side:
 * omg_free_space_kernel -- geometry of the FFT level + the kernel of 1/r in separable form, K(i, j, k) =
 *   sum_g w[g] kx[g][|i|] ky[g][|j|] kz[g][|k|] (n_range + 1 entries per Gaussian; what createKernel builds before its
 *   FFT), transformed once with cuFFT; also allots one Dirichlet plane to every physical-boundary face of every level.
 * omg_free_space_solve -- rhs of the FFT level (all ranks: all-reduce) * rhs_fac -> zero-padded cuFFT convolution ->
 *   potential on the (n+2)^3 grid * dr^3 -> the boundary planes of all levels by bilinear interpolation -> phi of the
 *   FFT level.  The caller then restricts / prolongs phi and cycles (omg_restrict_lvl, omg_prolong,
 *   omg_fill_ghost_cells_lvl, omg_fas_fmg / omg_fas_vcycle) exactly as :183-212 does. */
int omg_free_space_kernel(omg_handle* h, int fft_lvl, const double r_min[3], int n_gauss, int n_range, const double* w,
                          const double* kx, const double* ky, const double* kz);
int omg_free_space_solve(omg_handle* h, double rhs_fac);
/* host-side convenience: builds that separable kernel (order-8 interpolating scaling functions, 89-term Gaussian
 * quadrature of 1/r: createKernel('F', n_kernel(1), n_kernel(2), n_kernel(3), dr(1), dr(2), dr(3), 8, ...),
 * poisson_3d_fft/build_kernel.f90:55-199, 884-1050) and hands it to omg_free_space_kernel */
int omg_free_space_setup(omg_handle* h, int fft_lvl, const double r_min[3], const int n_kernel[3], const double dr[3]);
/* the stored boundary plane of a physical-boundary face (what mg_phi_bc_store keeps in the ghost cells of mg_irhs,
 * src/m_ghost_cells.f90:66-117): nc*nc values, first tangential index fastest; *bc_type receives its type */
int omg_get_bc_plane(omg_handle* h, int box_id, int nb, int* bc_type, double* plane);
/* repeat n V-cycles back to back without host round trips (benchmark loop of
 * tests/test_performance.f90:68-76); residuals[n] receives max_res per cycle when non-null */
int omg_fas_vcycles(omg_handle* h, int n, double* residuals);

/* One building block of the cycle on its own, for per-kernel parity tests: what = "smooth_boxes" (lvl,
 * arg = n_cycle; src/m_multigrid.f90:404-424), "update_coarse" (lvl; :347-384), "correct_children" (lvl;
 * :387-402 plus the ghost fill that follows), "coarse_solve" (:202-208), "max_residual_lvl" (lvl; :296-311,
 * result in *out). */
int omg_debug_step(omg_handle* h, const char* what, int lvl, int arg, double* out);
/* Self-test of the branch-free fp64 division the variable-coefficient kernels use (omg_device.cuh: ddiv_batch):
 * out[i] = a[i] / b[i] on the device, batched = 1 through ddiv_batch<6>, 0 through the compiler's `/`; host arrays. */
int omg_debug_ddiv(int device, const double* a, const double* b, double* out, int64_t n, int batched);
/* omg_debug_ddiv without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and result as omg_debug_ddiv (host arrays); the device staging that omg_debug_ddiv allocates lives in a
 * file-scope __device__ array of 3 x OMG_DDIV_STATIC_MAX doubles (6 MiB per GPU that loads the library, held until exit).
 * n > OMG_DDIV_STATIC_MAX is rejected (return 2, message in err); 1 = CUDA error, 3 = bad argument.  Calls no cudaMalloc,
 * cudaFree or allocating library routine.  All calls share the array, which keeps the last call's contents (nothing in it
 * has to start at zero): calls that overlap on one GPU race -- serialise them.  err (may be NULL) receives a message of at
 * most err_cap bytes. */
#define OMG_DDIV_STATIC_MAX 262144
int omg_debug_ddiv_static(int device, const double* a, const double* b, double* out, int64_t n, int batched, char* err, int err_cap);

/* ---- multi-GPU (replaces mg_comm_init + sort_and_transfer_buffers, src/m_communication.f90) */
int omg_nccl_unique_id(char out[128]);
int omg_comm_init(omg_handle* h, const char uid[128], int rank, int nranks);
/* Emulated ranks inside one process (one host thread per rank; parity tests on a single GPU): the handles
 * exchange through device copies ordered by CUDA events.  handles[r] must hold the tree of rank r. */
int omg_comm_init_local(omg_handle* const* handles, int nranks);
/* The static exchange plan, computable without a device: keys (2 * key + kind) of the items rank `me` sends to /
 * receives from `peer` in phase 0 (ghost cells, key 6*id + face, src/m_ghost_cells.f90:374-379), 1 (restriction,
 * key 8*parent + child, src/m_restrict.f90:158-160) or 2 (prolongation, key child id, src/m_prolong.f90:116-118)
 * on level lvl, in message order.  At most cap keys are written; the counts are always returned. */
int omg_exchange_plan(const omg_tree_desc* t, int me, int phase, int lvl, int peer, int64_t* send_keys,
                      int64_t* recv_keys, int cap, int* n_send, int* n_recv);

/* ---- diagnostics ----------------------------------------------------------- */
int omg_sync(omg_handle* h);                                          /* cudaStreamSynchronize */
int omg_timers(omg_handle* h, int* n, const char** names, double* seconds);  /* mg_timers_show */
int64_t omg_stat(omg_handle* h, const char* key);                     /* "launches", "bytes_h2d", ... */
/* Tuning / diagnostic switches (defaults in brackets); none changes results:
 *   "use_graph" [1] CUDA graph per cycle program, "graph_multi" [1] also with several ranks, "fast_kernels" [1] TMA-staged
 *   kernels where they apply, "tail_cells" [4096] levels up to this many cells run in the single-CTA tail, "tail_smem" [1]
 *   tail levels staged in shared memory, "tail_simple" [2: per-cell fast path of the tail for all operators and boundary kinds; 1 = Laplacian /
 *   Helmholtz between constant boundaries only; 0 = general functions], "warp_coarse" [1] coarse solve of a single 2^3 box by
 *   one warp with the cells in registers, "overlap" [1] halo exchange overlapped with interior boxes, "p2p" / "fused" [1] peer-mapped puts / issued
 *   from inside the smoother, "fused_prolong" [1] ... and from inside the prolongation kernel that precedes it, "put_fence" [2] one system-scope fence per
 *   CTA before its puts are signalled (1 = one per thread), "bnd_first_boxes", "pdl" [0] programmatic dependent launch (single rank; measured: no gain),
 *   "ring" [1] persistent ring smoother on big levels (2 = every level it serves, 0 = one tile per CTA),
 *   "store_fine_residual" [1] (0: the level's own residual, mg_ires, is not written by the cycle -- the one switch that
 *   changes what the host can read back: only for device-resident hosts that never read mg_ires),
 *   "profile" [0] per-kernel-class CUDA-event timing (omg_last_ms), "trace" [0] per-step timeline (omg_trace). */
int omg_set_option(omg_handle* h, const char* key, int value);
/* Per-step timeline (option "trace" = 1: every step of the cycle programs timed with its own CUDA-event pair, no graph,
 * in-kernel exchange left on): rec[i] = {step kind, level, a, b}, ms[i]; at most cap records are written, *n = number
 * recorded since the option was set or the last call; omg_step_name names a step kind. */
int omg_trace(omg_handle* h, int cap, int* rec, double* ms, int* n);
const char* omg_step_name(int kind);
/* CUDA-event timing of the last omg_fas_vcycles call (ms), and of one kernel class inside it */
double omg_last_ms(omg_handle* h, const char* what);
void* omg_stream(omg_handle* h);                                       /* cudaStream_t used by the handle */

#ifdef __cplusplus
}
#endif
#endif /* OMG_GPU_H */
