// omg_core.cu -- handle, tree ingestion, cycle programs and the C ABI (include/omg_gpu.h).
//
// The cycle (mg_fas_vcycle / mg_fas_fmg, src/m_multigrid.f90:84-243) is compiled once per call
// signature into a "program": a flat list of level operations.  Operations
// on the small coarse levels are grouped into op lists run by the single-CTA tail interpreter
// (one launch for the whole coarse end of the V, including the data-dependent coarse solve);
// operations on big levels become one batched kernel launch over all boxes of the level.  A
// program is replayed through a CUDA graph.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/omg_gpu.h"
#include "omg_comm.h"
#include "omg_fast.h"
#include "omg_fs.h"
#include "omg_kernels.h"
#include "omg_plan.h"

namespace {

constexpr int IPHI = 0, IRHS = 1, IOLD = 2, IRES = 3;

struct OmgError { std::string msg; };
[[noreturn]] void fail(const std::string& m) { throw OmgError{m}; }
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) fail(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

template <class T> struct DevBuf {
  T* p = nullptr; size_t n = 0;
  void alloc(size_t count) { free(); n = count; if (count) CK(cudaMalloc(&p, count * sizeof(T))); }
  void upload(const std::vector<T>& h, cudaStream_t s) {
    if (h.size() != n || (!p && n)) alloc(h.size());
    if (n) CK(cudaMemcpyAsync(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void free() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

enum StepKind : int {
  S_SMOOTH, S_RESID, S_RESTRICT, S_CRHS, S_OLDPL, S_CORR_PROLONG, S_RB_FILL, S_COPY_OLD, S_COARSE_SOLVE,
  S_TAIL, S_ZERO_PHI, S_SUBMEAN, S_CLEAR_SLOT, S_COPY_SLOT, S_HALO, S_XRESTRICT, S_XPROLONG, S_ALLREDUCE_MAX,
  S_RESID_RESTRICT, S_HALO_FORK, S_HALO_JOIN, S_PUT_WAIT
};
struct Step { int kind, l, a, b, c; };

// One communication phase of one level: per-peer segments (in items) and the pack / unpack descriptor lists
struct PhasePlan {
  struct Seg { int peer, send_off, send_n, recv_off, recv_n; };
  std::vector<Seg> seg;
  int n_send = 0, n_recv = 0;
  bool any = false;                     // some rank has traffic in this phase on this level (same on all ranks)
  std::vector<int> h_sdesc, h_rdesc;
  std::vector<int> h_dst;               // ghost phase as peer-mapped puts: per send item {receiver rank, plane in its pool}
  DevBuf<int> d_sdesc, d_rdesc, d_dst;
  std::vector<int> peers;               // ranks I exchange with in this phase
  DevBuf<double*> d_peer_base;          // [n_cpu] the peers' plane pools mapped on this device
  std::vector<long long> h_peer_stride;  // [n_cpu] distance between the two copies of each rank's pool (doubles)
  DevBuf<long long> d_peer_stride;
  bool shared = false;
  DevBuf<double> d_send, d_recv;
  void free() { d_sdesc.free(); d_rdesc.free(); d_dst.free(); d_peer_base.free(); d_peer_stride.free(); d_send.free(); d_recv.free(); }
};

struct LevelHost {
  int lvl = 0, nc = 0, nb = 0;
  PhasePlan xg, xr, xp;                 // ghost (onto this level), restrict (from this level), prolong (onto this level)
  bool tail = false;
  std::vector<int> ids;                 // owned boxes (global ids) in lvls(lvl)%my_ids order
  std::vector<int> h_fkind, h_farg, h_ftype, h_fkind_old, h_farg_old, h_parent, h_coff, h_parents, h_leaves,
      h_rb, h_oldpl;
  std::vector<double> h_bcpl;
  int n_gpl = 0, n_gpl_old = 0, n_bcpl = 0;
  bool meta_dirty = true;
  DevBuf<int> d_fkind, d_farg, d_ftype, d_fkind_old, d_farg_old, d_parent, d_coff, d_parents, d_leaves, d_rb, d_oldpl;
  DevBuf<double> d_gpl, d_gpl_old, d_bcpl;
  DevBuf<double> d_var[OMG_MAXV], d_faces[OMG_MAXV];
  LevelDev dev;
  // boxes with a same-level neighbour on another rank (they send and receive ghost faces) and the rest: the halo
  // exchange of a colour half-sweep overlaps the half-sweep of the interior boxes
  std::vector<int> h_bnd, h_int;
  DevBuf<int> d_bnd, d_int;
  // in-kernel exchange (PutDev): launch order = partition-face boxes, then the rest; per partition-face box the faces it
  // sends {receiver's face, receiver rank, plane in its pool}
  bool rb_remote_any = false;           // some rank receives interpolated coarse faces on this level (replicated-tree fact)
  std::vector<int> h_order, h_put_ptr, h_put_item, h_launch, h_launch_ring;   // h_launch_ring: interior boxes first (ring kernel)
  DevBuf<int> d_order, d_put_ptr, d_put_item, d_launch, d_launch_ring;
  DevBuf<unsigned int> d_done;
  bool all_fast = false;                // every face of every box has a kind the fast kernels serve (FK_SAME, FK_PHYS, FK_PLANE)
};

struct Program { std::vector<Step> steps; std::vector<TailOp> ops; DevBuf<TailOp> d_ops; cudaGraphExec_t graph = nullptr;
                 int64_t launches_per_run = 0, fast_per_run = 0, general_per_run = 0, fastc_per_run = 0, generalc_per_run = 0; };

}  // namespace

struct omg_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;          // halo exchanges that overlap interior smoothing
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  std::string err;
  // tree
  bool have_tree = false;
  int n_boxes = 0, box_size = 0, lowest = 0, highest = 0, first_normal = 0, my_rank = 0, n_cpu = 1, n_extra = 0;
  int periodic[3] = {0, 0, 0};
  int nvar = 4;
  std::vector<int> rank, blvl, ix, parent, children, neighbors, loc;
  std::vector<std::vector<int>> lvl_ids;
  std::vector<LevelHost> levels;
  DevBuf<LevelDev> d_levels;
  // parameters
  omg_params params{};
  bool have_params = false, params_dirty = true;
  int bc_type[OMG_MAXV][6];
  double bc_value[OMG_MAXV][6];
  ParamsDev pdev{};
  DevBuf<double> d_scal;
  DevBuf<double> d_partial;
  // programs
  std::map<std::string, Program> programs;
  bool gz[OMG_NLVL];
  // options
  int use_graph = 1, fast_kernels = 1, tail_cells = 4096, graph_multi = 1, tail_smem = 1, overlap = 1, tail_simple = 2, pdl = 0, ring = 1, store_fine_residual = 1, ring_order = 0, ring_fused = 0;
  int warp_coarse = 1;    // option "warp_coarse": the coarse solve of a single 2^3 box by one warp in registers (0: interpreter operations)
  int coeff_math = 0;     // option "coeff_math": 1 = tolerance mode of the coefficient kernels (omg_device.cuh recip6), default bitwise
  // cluster tail (k_tail_cluster): CTAs per cluster (0 = single-CTA tail only) and the level size it takes
  int tail_cluster = 0, tail_cluster_cells = 262144, tail_cluster_always = 0;   // off: measured slower (profiles/r02j_summary.txt)
  // staging
  double* h_stage[2] = {nullptr, nullptr};
  double* d_stage[2] = {nullptr, nullptr};
  int* d_boxidx[2] = {nullptr, nullptr};
  int* h_boxidx[2] = {nullptr, nullptr};
  size_t stage_doubles = 0; int stage_boxes = 0;
  cudaEvent_t stage_ev[2] = {nullptr, nullptr};
  // stats / timers
  int64_t launches = 0, bytes_h2d = 0, bytes_d2h = 0, graph_launches = 0, exchanges = 0, fused_smooths = 0, fused_prolongs = 0, puts = 0;
  // level kernels launched through the TMA-staged path (omg_fast.cu) / through the general one-CTA-per-box path
  // (omg_kernels.cu); the single-CTA tail and the pack / exchange kernels count in neither
  int64_t fast_launches = 0, general_launches = 0;
  int64_t fast_cells = 0, general_cells = 0;   // the same, weighted by the cells of the level each launch ran on
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double last_ms = 0.0;
  omg_comm* comm = nullptr;
  int bnd_first_boxes = 8192;               // levels up to this many boxes launch their partition-face boxes first
  int fused = 1;                            // in-kernel exchange in the smoother where fused_level() allows (option "fused")
  int put_fence = 2;                        // in-kernel puts: 2 (default) = one system-scope fence per CTA after the CTA barrier, 1 = every thread fences its own remote stores (option "put_fence"; 2 GPUs: 4.737 vs 4.792 ms per cycle, profiles/r02zg_*)
  int fused_prolong = 1;                    // ... and in the prolongation kernel that precedes such a smoother (option "fused_prolong")
  int p2p = 1;                              // ghost faces as peer-mapped puts where the plane pools are shared (option "p2p")
  DevBuf<unsigned long long> d_epoch;       // [OMG_NLVL] completed ghost exchanges per level (selects the plane-pool copy)
  // per-step CUDA-event profile of the finest level (option "profile"): kind -> (events, count)
  int profile = 0;
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof_events;   // first = kind * 2 + (finest level ? 1 : 0)
  // the reference's timers (mg_add_timers, src/m_multigrid.f90:72-81), seconds; totals always, phases in "profile" mode
  double timer_acc[7] = {0, 0, 0, 0, 0, 0, 0};
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> total_events;   // (timer index, events) not yet read
  std::map<int, std::pair<double, int64_t>> prof_acc;   // step kind -> (ms, launches)
  // option "trace": every step of every program run timed with its own event pair, cycle programs unchanged (the in-kernel
  // exchange stays on, unlike "profile"); read with omg_trace
  int trace = 0;
  struct TraceRec { int kind, l, a, b; cudaEvent_t e0, e1; };
  std::vector<TraceRec> trace_events;

  // free-space front end (omg_fs.h): plan of the FFT level, box origins in its grid, the boundary faces of every level
  omgfs::Plan* fs_plan = nullptr;
  int fs_lvl = 0;
  omgfs::Geom fs_geom{};
  DevBuf<int> fs_origin;
  std::vector<DevBuf<omgfs::Face>> fs_faces;     // per level
  std::vector<int> fs_nfaces;

  LevelHost& L(int lvl) { return levels[lvl - lowest]; }
};

namespace {

void invalidate_programs(omg_handle* h) {
  for (auto& kv : h->programs) { if (kv.second.graph) cudaGraphExecDestroy(kv.second.graph); kv.second.d_ops.free(); }
  h->programs.clear();
}

void free_levels(omg_handle* h) {
  for (LevelHost& L : h->levels) {
    L.d_fkind.free(); L.d_farg.free(); L.d_ftype.free(); L.d_fkind_old.free(); L.d_farg_old.free();
    L.d_parent.free(); L.d_coff.free(); L.d_parents.free(); L.d_leaves.free(); L.d_rb.free(); L.d_oldpl.free();
    L.d_gpl.free(); L.d_gpl_old.free(); L.d_bcpl.free(); L.d_bnd.free(); L.d_int.free(); L.d_order.free(); L.d_launch.free(); L.d_launch_ring.free(); L.d_put_ptr.free(); L.d_put_item.free(); L.d_done.free();
    for (int v = 0; v < OMG_MAXV; ++v) { L.d_var[v].free(); L.d_faces[v].free(); }
    L.xg.free(); L.xr.free(); L.xp.free();
  }
  h->levels.clear();
  h->d_levels.free();
  if (h->fs_plan) { omgfs::destroy(h->fs_plan); h->fs_plan = nullptr; }
  h->fs_origin.free();
  for (auto& f : h->fs_faces) f.free();
  h->fs_faces.clear(); h->fs_nfaces.clear();
  invalidate_programs(h);
}

inline bool has_children(const omg_handle* h, int id) { return h->children[(size_t)(id - 1) * 8] != 0; }

// tail group: the coarsest levels, while they are small and live on one rank (the reference keeps
// everything below first_normal_lvl on one rank, src/m_load_balance.f90:79-81)
// the cluster tail serves GSRB on even box sizes (the other smoothers synchronise inside a CTA)
bool cluster_tail_ok(const omg_handle* h) {
  if (h->tail_cluster < 2 || h->params.smoother_type != 2) return false;
  for (const LevelHost& L : h->levels) if (L.nc & 1) return false;
  return true;
}
void compute_tail(omg_handle* h) {
  bool still = true; int owner = -1;
  for (int lvl = h->lowest; lvl <= h->highest; ++lvl) {
    LevelHost& L = h->L(lvl);
    const std::vector<int>& ids = h->lvl_ids[lvl - h->lowest];
    bool one_rank = true;
    for (int id : ids) { if (owner < 0) owner = h->rank[id - 1]; if (h->rank[id - 1] != owner) one_rank = false; }
    const long cells = (long)ids.size() * L.nc * L.nc * L.nc;
    if (lvl == h->lowest) {
      if (!one_rank) fail("Multiple CPUs for coarse grid (not implemented yet)");   // m_multigrid.f90:197-200
      L.tail = true;
    } else {
      still = still && one_rank && cells <= std::max(h->tail_cells, cluster_tail_ok(h) ? h->tail_cluster_cells : 0) && lvl < h->highest;
      L.tail = still;
    }
  }
}

// ---- tree ingestion ------------------------------------------------------------------------
void set_tree(omg_handle* h, const omg_tree_desc* t) {
  CK(cudaSetDevice(h->device));
  free_levels(h);
  h->n_boxes = t->n_boxes; h->box_size = t->box_size; h->lowest = t->lowest_lvl; h->highest = t->highest_lvl;
  h->first_normal = t->first_normal_lvl; h->my_rank = t->my_rank; h->n_cpu = t->n_cpu; h->n_extra = t->n_extra_vars;
  for (int d = 0; d < 3; ++d) h->periodic[d] = t->periodic[d];
  h->nvar = 4 + h->n_extra;
  if (h->nvar > OMG_MAXV) fail("omg_set_tree: too many variables");
  if (h->lowest < OMG_LVL_LO || h->highest > OMG_LVL_HI || h->lowest > h->highest) fail("omg_set_tree: bad level range");
  const size_t n = (size_t)t->n_boxes;
  h->rank.assign(t->box_rank, t->box_rank + n);
  h->blvl.assign(t->box_lvl, t->box_lvl + n);
  h->ix.assign(t->box_ix, t->box_ix + 3 * n);
  h->parent.assign(t->box_parent, t->box_parent + n);
  h->children.assign(t->box_children, t->box_children + 8 * n);
  h->neighbors.assign(t->box_neighbors, t->box_neighbors + 6 * n);
  h->loc.assign(n + 1, -1);
  const int nl = h->highest - h->lowest + 1;
  h->levels.assign(nl, LevelHost());
  h->lvl_ids.assign(nl, std::vector<int>());
  const int me = h->my_rank;

  for (int lvl = h->lowest; lvl <= h->highest; ++lvl) {
    LevelHost& L = h->L(lvl);
    L.lvl = lvl; L.nc = t->box_size_lvl[lvl - OMG_LVL_LO];
    if (L.nc < 1 || L.nc > 1023) fail("omg_set_tree: unsupported box size");
    const int a = t->lvl_ptr[lvl - OMG_LVL_LO], b = t->lvl_ptr[lvl - OMG_LVL_LO + 1];
    h->lvl_ids[lvl - h->lowest].assign(t->lvl_ids + a, t->lvl_ids + b);
    for (int q = a; q < b; ++q) {
      const int id = t->lvl_ids[q];
      if (id < 1 || id > t->n_boxes) fail("omg_set_tree: box id out of range");
      if (h->rank[id - 1] == me) { h->loc[id] = (int)L.ids.size(); L.ids.push_back(id); }
    }
    L.nb = (int)L.ids.size();
  }

  compute_tail(h);

  // which (phase, level) pairs carry traffic on ANY rank: decided from the replicated tree, so every rank
  // emits the same exchange steps
  if (h->n_cpu > 1) {
    for (int id = 1; id <= h->n_boxes; ++id) {
      const int lvl = h->blvl[id - 1];
      if (lvl < h->lowest || lvl > h->highest) continue;
      LevelHost& L = h->L(lvl);
      const int r = h->rank[id - 1], p = h->parent[id - 1];
      if (p > 0 && h->rank[p - 1] != r) { L.xr.any = true; L.xp.any = true; }
      for (int f = 0; f < 6; ++f) {
        const int nbid = h->neighbors[(size_t)(id - 1) * 6 + f];
        if (nbid > 0) { if (h->rank[nbid - 1] != r) L.xg.any = true; }
        else if (nbid == 0 && p > 0) {
          L.rb_remote_any = true;       // any refinement boundary on the level rules the in-kernel exchange out (its
                                        // planes are refilled by a separate kernel after every exchange)
          const int cnb = h->neighbors[(size_t)(p - 1) * 6 + f];
          if (cnb > 0 && h->rank[cnb - 1] != r) { L.xg.any = true; L.rb_remote_any = true; }
        }
      }
    }
  }
  auto coff_of = [&](int id, int lvl) {
    if (lvl <= h->first_normal) return 0;
    int off[3];
    for (int d = 0; d < 3; ++d) off[d] = ((h->ix[(size_t)(id - 1) * 3 + d] - 1) & 1) * (h->box_size >> 1);
    return off[0] | (off[1] << 10) | (off[2] << 20);
  };

  for (int lvl = h->lowest; lvl <= h->highest; ++lvl) {
    LevelHost& L = h->L(lvl);
    const int nb = L.nb, nc = L.nc;
    // ghost-phase plan: planes received from peer ranks go straight into the plane pool, one contiguous
    // run per peer (same-level faces first, then interpolated coarse faces), see omg_plan.h
    std::map<std::pair<int, int>, int> remote_plane;
    std::map<int, std::vector<int>> put_by_box;          // sender's local box -> {receiver face, rank, plane}*
    if (L.xg.any) {
      omgp::LevelPlan pl = omgp::build_plan(*t, me, omgp::PH_GHOST, lvl);
      int soff = 0, roff = 0;
      // refinement-boundary faces per rank (each owns one plane in its pool, after the planes it receives)
      std::vector<long long> n_rb_of(h->n_cpu, 0);
      for (int id : h->lvl_ids[lvl - h->lowest])
        for (int f = 0; f < 6; ++f) if (h->neighbors[(size_t)(id - 1) * 6 + f] == 0) n_rb_of[h->rank[id - 1]]++;
      L.xg.h_peer_stride.assign(h->n_cpu, 1);
      for (const omgp::PeerPlan& pp : pl.peers) {
        L.xg.seg.push_back({pp.peer, soff, (int)pp.send.size(), roff, (int)pp.recv.size()});
        for (const omgp::Item& it : pp.recv) remote_plane[{it.box, it.face}] = roff++;
        // where my items land in the receiver's plane pool: its recv planes are numbered peer by peer in its own plan
        int rbase = -1;
        {
          omgp::LevelPlan theirs = omgp::build_plan(*t, pp.peer, omgp::PH_GHOST, lvl);
          int ro = 0;
          for (const omgp::PeerPlan& q : theirs.peers) { if (q.peer == me) { rbase = ro; if (q.recv.size() != pp.send.size()) fail("omg_set_tree: exchange plans of two ranks disagree"); } ro += (int)q.recv.size(); }
          L.xg.h_peer_stride[pp.peer] = std::max<long long>((ro + n_rb_of[pp.peer]) * (long long)nc * nc, 1);   // = its d_gpl.n / 2
        }
        if (!pp.send.empty() && rbase < 0) fail("omg_set_tree: exchange plans of two ranks disagree");
        int k = 0;
        for (const omgp::Item& it : pp.send) {
          L.xg.h_sdesc.push_back(it.kind); L.xg.h_sdesc.push_back(h->loc[it.src]); L.xg.h_sdesc.push_back(it.face);
          L.xg.h_sdesc.push_back(it.kind == omgp::IT_RB ? coff_of(it.box, lvl) : 0);
          L.xg.h_dst.push_back(pp.peer); L.xg.h_dst.push_back(rbase + k);
          if (it.kind == omgp::IT_SAME) { std::vector<int>& v = put_by_box[h->loc[it.src]]; v.push_back(it.face); v.push_back(pp.peer); v.push_back(rbase + k); }
          ++k;
          ++soff;
        }
        L.xg.peers.push_back(pp.peer);
      }
      L.xg.n_send = soff; L.xg.n_recv = roff;
      L.n_gpl = roff;
    }
    L.h_fkind.assign((size_t)nb * 6, FK_NONE); L.h_farg.assign((size_t)nb * 6, -1); L.h_ftype.assign((size_t)nb * 6, 0);
    L.h_fkind_old.assign((size_t)nb * 6, FK_NONE); L.h_farg_old.assign((size_t)nb * 6, -1);
    L.h_parent.assign(nb, -1); L.h_coff.assign(nb, 0);
    L.all_fast = true;                 // cleared below if a face keeps FK_NONE
    for (int b = 0; b < nb; ++b) {
      const int id = L.ids[b];
      const bool is_parent = has_children(h, id);
      (is_parent ? L.h_parents : L.h_leaves).push_back(b);
      bool on_partition_face = false;
      for (int f = 0; f < 6; ++f) {
        const int nbid = h->neighbors[(size_t)(id - 1) * 6 + f];
        if (nbid > 0 && h->rank[nbid - 1] != me) on_partition_face = true;
        if (nbid == 0) on_partition_face = true;    // refinement-boundary planes live in the pool copy selected by the epoch,
                                                    // which flips while the "interior" boxes are being swept
      }
      (on_partition_face ? L.h_bnd : L.h_int).push_back(b);
      const int p = h->parent[id - 1];
      if (p > 0 && h->rank[p - 1] == me) L.h_parent[b] = h->loc[p];
      L.h_coff[b] = coff_of(id, lvl);
      for (int f = 0; f < 6; ++f) {
        const int nbid = h->neighbors[(size_t)(id - 1) * 6 + f];
        const size_t q = (size_t)b * 6 + f;
        if (nbid > 0) {
          if (h->rank[nbid - 1] == me) {
            L.h_fkind[q] = FK_SAME; L.h_farg[q] = h->loc[nbid];
            if (!is_parent || has_children(h, nbid)) { L.h_fkind_old[q] = FK_SAME; L.h_farg_old[q] = h->loc[nbid]; }
            else { L.h_fkind_old[q] = FK_PLANE; L.h_farg_old[q] = L.n_gpl_old;
                   L.h_oldpl.push_back(b); L.h_oldpl.push_back(f); L.h_oldpl.push_back(L.n_gpl_old++); }
          } else {
            L.h_fkind[q] = FK_PLANE; L.h_farg[q] = remote_plane.at({id, f});   // filled by the halo exchange
            if (is_parent) { L.h_fkind_old[q] = FK_PLANE; L.h_farg_old[q] = L.n_gpl_old;
                             L.h_oldpl.push_back(b); L.h_oldpl.push_back(f); L.h_oldpl.push_back(L.n_gpl_old++); }
          }
        } else if (nbid == 0) {
          if (p <= 0) fail("omg_set_tree: refinement boundary on a box without parent");
          const int cnb = h->neighbors[(size_t)(p - 1) * 6 + f];
          if (cnb <= 0) fail("omg_set_tree: refinement boundary without coarse neighbour (tree not 2:1 balanced)");
          const int cb = (h->rank[cnb - 1] == me) ? h->loc[cnb] : -(2 + remote_plane.at({id, f}));
          L.h_fkind[q] = FK_PLANE; L.h_farg[q] = L.n_gpl;
          L.h_rb.push_back(b); L.h_rb.push_back(f); L.h_rb.push_back(cb); L.h_rb.push_back(L.n_gpl++);
        } else {
          L.h_fkind[q] = FK_PHYS; L.h_farg[q] = -1; L.h_ftype[q] = (nbid == -1) ? 0 : nbid;
          L.h_fkind_old[q] = FK_PHYS; L.h_farg_old[q] = -1;
        }
        if (L.h_fkind[q] == FK_NONE) L.all_fast = false;
      }
    }
    L.h_order.assign(nb, -1);                          // box -> index among the partition-face boxes
    for (size_t j = 0; j < L.h_bnd.size(); ++j) L.h_order[L.h_bnd[j]] = (int)j;
    // launch order with the partition-face boxes first: on small levels their faces then leave early in the kernel
    // (on big levels the extra indirection at every CTA start costs more than it hides, measured)
    L.h_launch = L.h_bnd;
    L.h_launch.insert(L.h_launch.end(), L.h_int.begin(), L.h_int.end());
    L.h_launch_ring = L.h_int;
    L.h_launch_ring.insert(L.h_launch_ring.end(), L.h_bnd.begin(), L.h_bnd.end());
    L.h_put_ptr.assign(1, 0);
    for (int b : L.h_bnd) {
      auto it = put_by_box.find(b);
      if (it != put_by_box.end()) L.h_put_item.insert(L.h_put_item.end(), it->second.begin(), it->second.end());
      L.h_put_ptr.push_back((int)L.h_put_item.size() / 3);
    }
    // restrict (children on this level -> parents on lvl-1) and prolong (parents on lvl-1 -> children here)
    if (L.xr.any) {
      omgp::LevelPlan pl = omgp::build_plan(*t, me, omgp::PH_RESTRICT, lvl);
      int soff = 0, roff = 0;
      for (const omgp::PeerPlan& pp : pl.peers) {
        L.xr.seg.push_back({pp.peer, soff, (int)pp.send.size(), roff, (int)pp.recv.size()});
        for (const omgp::Item& it : pp.send) { L.xr.h_sdesc.push_back(h->loc[it.src]); ++soff; }
        for (const omgp::Item& it : pp.recv) { L.xr.h_rdesc.push_back(h->loc[it.box]); L.xr.h_rdesc.push_back(coff_of(it.src, lvl)); ++roff; }
      }
      L.xr.n_send = soff; L.xr.n_recv = roff;
      const size_t item = (size_t)2 * (nc / 2) * (nc / 2) * (nc / 2);
      L.xr.d_send.alloc(soff * item); L.xr.d_recv.alloc(roff * item);
    }
    if (L.xp.any) {
      omgp::LevelPlan pl = omgp::build_plan(*t, me, omgp::PH_PROLONG, lvl);
      int soff = 0, roff = 0;
      for (const omgp::PeerPlan& pp : pl.peers) {
        L.xp.seg.push_back({pp.peer, soff, (int)pp.send.size(), roff, (int)pp.recv.size()});
        for (const omgp::Item& it : pp.send) { L.xp.h_sdesc.push_back(h->loc[it.src]); L.xp.h_sdesc.push_back(coff_of(it.box, lvl)); ++soff; }
        for (const omgp::Item& it : pp.recv) L.h_parent[h->loc[it.box]] = -(2 + roff++);
      }
      L.xp.n_send = soff; L.xp.n_recv = roff;
      const size_t item = (size_t)(nc / 2 + 2) * (nc / 2 + 2) * (nc / 2 + 2);
      L.xp.d_send.alloc(soff * item); L.xp.d_recv.alloc(roff * item);
    }
    if (L.xg.any) L.xg.d_send.alloc((size_t)L.xg.n_send * nc * nc);
    // storage: mg_allocate_storage zero-fills every variable (src/m_allocate_storage.f90:72-76)
    const int hx = (nc + 1) / 2;
    const size_t bs = (size_t)2 * nc * nc * hx;
    for (int v = 0; v < h->nvar; ++v) {
      L.d_var[v].alloc((size_t)nb * bs);
      if (nb) CK(cudaMemsetAsync(L.d_var[v].p, 0, (size_t)nb * bs * sizeof(double), h->stream));
      if (v >= 4) {
        L.d_faces[v].alloc((size_t)nb * 6 * nc * nc);
        if (nb) CK(cudaMemsetAsync(L.d_faces[v].p, 0, (size_t)nb * 6 * nc * nc * sizeof(double), h->stream));
      }
    }
    // distributed levels keep two copies of the plane pool (peer-mapped puts fill one while the other is read,
    // omg_device.cuh); every rank allocates it, even with no planes of its own, so that the mapping is collective
    L.d_gpl.alloc(L.xg.any ? std::max<size_t>((size_t)2 * L.n_gpl * nc * nc, 2) : (size_t)L.n_gpl * nc * nc);
    if (L.d_gpl.n) CK(cudaMemsetAsync(L.d_gpl.p, 0, L.d_gpl.n * sizeof(double), h->stream));
    L.d_gpl_old.alloc((size_t)L.n_gpl_old * nc * nc);
    if (L.n_gpl_old) CK(cudaMemsetAsync(L.d_gpl_old.p, 0, L.d_gpl_old.n * sizeof(double), h->stream));
    L.meta_dirty = true;

    LevelDev& D = L.dev;
    std::memset(&D, 0, sizeof(D));
    D.lvl = lvl; D.nc = nc; D.hx = hx; D.nb = nb; D.cs = nc * nc * hx; D.bs = 2 * D.cs;
    D.sh_nc = D.sh_hx = -1;
    if (nc >= 2 && (nc & (nc - 1)) == 0) { D.sh_nc = 0; while ((1 << D.sh_nc) < nc) ++D.sh_nc; D.sh_hx = D.sh_nc - 1; }
    for (int d = 0; d < 3; ++d) {
      const double dr = t->dr[(size_t)(lvl - OMG_LVL_LO) * 3 + d];
      D.dr[d] = dr; D.idr2[d] = 1 / (dr * dr);
    }
  }
  for (int l = 0; l < OMG_NLVL; ++l) h->gz[l] = false;
  h->have_tree = true;
  h->params_dirty = true;
  size_t maxleaves = 1;
  for (LevelHost& L : h->levels) maxleaves = std::max(maxleaves, L.h_leaves.size());
  h->d_partial.alloc(maxleaves);
}

// May the smoother of this level exchange its partition faces in-kernel?  Every input is the same on all ranks
// (replicated tree, parameters, options, and the collectively established pool mapping).
bool fused_level(const omg_handle* h, const LevelHost& L) {
  if (!h->comm || !h->p2p || !h->fused || !L.xg.any || !L.xg.shared || L.rb_remote_any || L.tail) return false;
  if (L.nc != 16 && L.nc != 32) return false;
  const int op = h->params.operator_type;
  if (op < 1 || op > 5 || h->params.smoother_type != 2 || !h->fast_kernels) return false;
  return h->n_cpu <= 9;                       // PutDev holds up to 8 peers
}

// upload metadata that changed (face tables, bc planes, parameters) and refresh the LevelDev copies
void sync_meta(omg_handle* h) {
  if (!h->have_tree) fail("no tree: call omg_set_tree first");
  if (!h->have_params) fail("no parameters: call omg_set_params first");
  bool any = h->params_dirty;
  for (LevelHost& L : h->levels) any = any || L.meta_dirty;
  if (!any) return;
  cudaStream_t s = h->stream;
  if (h->params_dirty) compute_tail(h);      // the tail group depends on the smoother (cluster tail: GSRB only)
  for (LevelHost& L : h->levels) {
    LevelDev& D = L.dev;
    if (L.meta_dirty) {
      L.d_fkind.upload(L.h_fkind, s); L.d_farg.upload(L.h_farg, s); L.d_ftype.upload(L.h_ftype, s);
      L.d_fkind_old.upload(L.h_fkind_old, s); L.d_farg_old.upload(L.h_farg_old, s);
      L.d_parent.upload(L.h_parent, s); L.d_coff.upload(L.h_coff, s);
      L.d_parents.upload(L.h_parents, s); L.d_leaves.upload(L.h_leaves, s);
      L.d_rb.upload(L.h_rb, s); L.d_oldpl.upload(L.h_oldpl, s);
      L.d_bnd.upload(L.h_bnd, s); L.d_int.upload(L.h_int, s);
      L.d_order.upload(L.h_order, s); L.d_launch.upload(L.h_launch, s); L.d_launch_ring.upload(L.h_launch_ring, s); L.d_put_ptr.upload(L.h_put_ptr, s); L.d_put_item.upload(L.h_put_item, s);
      if (!L.d_done.p) { L.d_done.alloc(1); CK(cudaMemsetAsync(L.d_done.p, 0, sizeof(unsigned int), s)); }
      L.d_bcpl.upload(L.h_bcpl, s);
      for (PhasePlan* X : {&L.xg, &L.xr, &L.xp}) { X->d_sdesc.upload(X->h_sdesc, s); X->d_rdesc.upload(X->h_rdesc, s); X->d_dst.upload(X->h_dst, s); X->d_peer_stride.upload(X->h_peer_stride, s); }
      if (L.xp.d_recv.n) CK(cudaMemsetAsync(L.xp.d_recv.p, 0, L.xp.d_recv.n * sizeof(double), s));
      L.meta_dirty = false;
    }
    if (h->params.smoother_type == 1 && L.nb && !L.d_faces[IPHI].p)   // mg_smoother_gs: ghost snapshot of phi
      L.d_faces[IPHI].alloc((size_t)L.nb * 6 * L.nc * L.nc);
    for (int v = 0; v < OMG_MAXV; ++v) { D.var[v] = L.d_var[v].p; D.faces[v] = L.d_faces[v].p; }
    D.fkind = L.d_fkind.p; D.farg = L.d_farg.p; D.ftype = L.d_ftype.p;
    D.fkind_old = L.d_fkind_old.p; D.farg_old = L.d_farg_old.p;
    D.gpl = L.d_gpl.p; D.gpl_old = L.d_gpl_old.p; D.bcpl = L.d_bcpl.p;
    const bool put = h->p2p && L.xg.shared;
    D.gpl_sel = put ? h->d_epoch.p + (L.lvl - OMG_LVL_LO) : nullptr;
    D.gpl_stride = (long long)(L.d_gpl.n / 2);
    std::memset(&D.put, 0, sizeof(D.put));
    if (fused_level(h, L)) {
      PutDev& U = D.put;
      U.n_bnd = (int)L.h_bnd.size(); U.me = h->my_rank; U.nranks = h->n_cpu; U.channel = L.lvl - OMG_LVL_LO;
      U.n_peers = (int)L.xg.peers.size();
      for (int q = 0; q < U.n_peers; ++q) U.peers[q] = L.xg.peers[q];
      U.slot_of = L.d_order.p; U.ptr = L.d_put_ptr.p; U.item = L.d_put_item.p; U.peer_base = L.xg.d_peer_base.p; U.peer_stride = L.xg.d_peer_stride.p;
      unsigned long long* const* pf = nullptr; unsigned long long* fl = nullptr; int* er = nullptr;
      try { omg_comm_p2p_ptrs(h->comm, &pf, &fl, &er); } catch (const std::exception& e) { fail(e.what()); }
      U.peer_flags = pf; U.flags = fl; U.err = er; U.timeout_ns = omg_comm_p2p_timeout_ns(h->comm);
      U.epoch = h->d_epoch.p + (L.lvl - OMG_LVL_LO); U.done = L.d_done.p;
    }
    // tail fast path (omg_ops.cuh simple_*): every face a physical boundary or a local same-level neighbour
    // Option "tail_simple": 2 (default) = all five operators, constant boundary values and stored boundary planes (bitwise on
    // a B200: tests/test_gpu_steps.py test_tail_fast_path_all_variants, the whole GPU suite since pass zb); 1 = Laplacian /
    // Helmholtz between constant boundaries only; 0 = the general per-cell functions.
    D.simple = (h->tail_simple && L.nb >= 1 && D.sh_nc >= 0) ? 1 : 0;
    const bool all_variants = h->tail_simple >= 2;
    if (!all_variants && h->params.operator_type != 1 && h->params.operator_type != 3) D.simple = 0;
    for (size_t q = 0; q < L.h_fkind.size() && D.simple; ++q) {
      const bool phi_ok = L.h_fkind[q] == FK_SAME || (L.h_fkind[q] == FK_PHYS && (all_variants || L.h_farg[q] < 0));
      const bool old_ok = L.h_fkind_old[q] == FK_SAME || (L.h_fkind_old[q] == FK_PHYS && (all_variants || L.h_farg_old[q] < 0));
      if (!phi_ok || !old_ok) D.simple = 0;
    }
    D.parent = L.d_parent.p; D.coff = L.d_coff.p; D.ppatch = L.xp.d_recv.p;
    D.parents_list = L.d_parents.p; D.n_parents = (int)L.h_parents.size();
    D.leaves_list = L.d_leaves.p; D.n_leaves = (int)L.h_leaves.size();
    D.rb_list = L.d_rb.p; D.n_rb = (int)L.h_rb.size() / 4;
    D.oldpl_list = L.d_oldpl.p; D.n_oldpl = (int)L.h_oldpl.size() / 3;
    const double sum = D.idr2[0] + D.idr2[1] + D.idr2[2];
    D.fac = (h->params.operator_type == OP_HELMH) ? 1.0 / (2 * sum + h->params.lambda) : 0.5 / sum;
  }
  std::vector<LevelDev> devs;
  for (LevelHost& L : h->levels) devs.push_back(L.dev);
  h->d_levels.upload(devs, s);
  h->pdev.op = h->params.operator_type; h->pdev.smoother = h->params.smoother_type; h->pdev.lambda = h->params.lambda;
  h->pdev.coeff_math = h->coeff_math; h->pdev.warp_coarse = h->warp_coarse;
  for (int f = 0; f < 6; ++f) { h->pdev.bc_type[f] = h->bc_type[IPHI][f]; h->pdev.bc_value[f] = h->bc_value[IPHI][f]; }
  if (!h->d_scal.p) { h->d_scal.alloc(SCAL_COUNT); CK(cudaMemsetAsync(h->d_scal.p, 0, SCAL_COUNT * sizeof(double), s)); }
  const double ra[2] = {h->params.residual_coarse_rel, h->params.residual_coarse_abs};
  CK(cudaMemcpyAsync(h->d_scal.p + SCAL_COARSE_REL, ra, sizeof(ra), cudaMemcpyHostToDevice, s));
  CK(cudaStreamSynchronize(s));   // host vectors / stack values above must outlive the copies
  h->params_dirty = false;
  invalidate_programs(h);
}

// ---- program builder: the reference's control flow, flattened ---------------------------------
struct Builder {
  omg_handle* h; Program& P; std::vector<TailOp> pend;
  bool multi() const { return h->n_cpu > 1; }
  int li(int lvl) const { return lvl - h->lowest; }
  void flush() {
    if (pend.empty()) return;
    P.steps.push_back({S_TAIL, 0, (int)P.ops.size(), (int)pend.size(), 0});
    P.ops.insert(P.ops.end(), pend.begin(), pend.end());
    pend.clear();
  }
  void emit(int lvl, int tail_type, int step_kind, int a = 0, int b = 0) {
    if (h->L(lvl).tail) pend.push_back({tail_type, li(lvl), a, b});
    else { flush(); P.steps.push_back({step_kind, lvl, a, b, 0}); }
  }
  void step(int kind, int lvl, int a = 0, int b = 0, int c = 0) { flush(); P.steps.push_back({kind, lvl, a, b, c}); }

  // mg_fill_ghost_cells_lvl(lvl, mg_iphi): only ghost values with remote inputs are materialised
  int halo_flag(int lvl) const { return (lvl > h->lowest && h->gz[lvl - 1 - OMG_LVL_LO]) ? 1 : 0; }
  void gf_local(int lvl) {
    if (!h->L(lvl).h_rb.empty()) emit(lvl, TOP_RB_FILL, S_RB_FILL, h->gz[lvl - 1 - OMG_LVL_LO] ? 1 : 0);
    h->gz[lvl - OMG_LVL_LO] = false;
  }
  void gf(int lvl) {
    if (h->L(lvl).xg.any) step(S_HALO, lvl, halo_flag(lvl));
    gf_local(lvl);
  }
  // smooth_boxes src/m_multigrid.f90:404-424.  On a distributed level a red-black half-sweep runs boundary-first:
  // boxes on partition faces, then [pack + NCCL exchange on the second stream || the interior boxes], then join
  // (a GSRB half-sweep reads only the other colour, so the two subsets are independent; the decision uses the
  // replicated tree only, so every rank emits the same steps).
  void smooth_boxes(int lvl, int n_cycle) {
    const int sub = (h->params.smoother_type == 2) ? 2 : 1;
    const bool fusedx = !h->profile && fused_level(h, h->L(lvl));
    const bool split = !fusedx && h->overlap && !h->profile && sub == 2 && h->L(lvl).xg.any && !h->L(lvl).tail;
    for (int n = 1; n <= n_cycle * sub; ++n) {
      if (fusedx) {
        // the partition-face boxes of this launch wait for the peers' previous faces, store their own into the peers'
        // pools and signal; consumers other than the next half-sweep need the explicit wait emitted after the loop
        step(S_SMOOTH, lvl, (n - 1) & 1, 3);
        h->gz[lvl - OMG_LVL_LO] = false;
        if (n == n_cycle * sub) step(S_PUT_WAIT, lvl);
      } else if (split) {
        step(S_SMOOTH, lvl, (n - 1) & 1, 1);
        step(S_HALO_FORK, lvl, halo_flag(lvl));
        step(S_SMOOTH, lvl, (n - 1) & 1, 2);
        step(S_HALO_JOIN, lvl);
        gf_local(lvl);
      } else { emit(lvl, TOP_SMOOTH, S_SMOOTH, (n - 1) & 1); gf(lvl); }
    }
  }
  // update_coarse :347-384
  void update_coarse(int lvl) {
    const bool fuse = !h->L(lvl).tail;
    if (fuse) step(S_RESID_RESTRICT, lvl);
    else {
      emit(lvl, TOP_RESID, S_RESID, -1);
      emit(lvl, TOP_RESTRICT, S_RESTRICT, IPHI, IRES);
    }
    if (h->L(lvl).xr.any) step(S_XRESTRICT, lvl, IPHI, IRES);
    gf(lvl - 1);
    emit(lvl - 1, TOP_CRHS, S_CRHS);
    if (!h->L(lvl - 1).h_oldpl.empty()) emit(lvl - 1, TOP_OLDPL, S_OLDPL);
  }
  // correct_children(lvl-1) :387-402, prolonging onto lvl, and the ghost-cell fill that follows it (:219-222, :132-136).
  // Where the next operation on the level is a GSRB half-sweep with in-kernel exchange (n_smooth > 0 of them follow), the
  // partition-face boxes of the prolongation kernel put their corrected faces into the peers' plane pools themselves and the
  // half-sweep's partition-face CTAs wait for the peers' signal: no stand-alone exchange (pack + put, signal + wait) in between.
  // Every input of the decision is replicated, so all ranks emit the same steps.
  void correct_children_onto(int lvl, int n_smooth) {
    if (h->L(lvl).xp.any) step(S_XPROLONG, lvl);
    const LevelHost& L = h->L(lvl);
    const bool put = h->fused_prolong && n_smooth > 0 && h->params.smoother_type == 2 && !h->profile && fused_level(h, L) &&
                     lvl > h->lowest && h->L(lvl - 1).nc == L.nc;
    if (put) { step(S_CORR_PROLONG, lvl, 3); gf_local(lvl); }
    else { emit(lvl, TOP_CORR_PROLONG, S_CORR_PROLONG); gf(lvl); }
  }
  void subtract_mean(int iv, int with_ghosts) { step(S_SUBMEAN, 0, iv, with_ghosts); }

  // mg_fas_vcycle :150-243
  void vcycle(bool has_highest, int highest_lvl, bool standalone, bool want_res, int res_slot) {
    if (h->params.subtract_mean && !has_highest) subtract_mean(IRHS, 0);
    const int min_lvl = h->lowest, max_lvl = has_highest ? highest_lvl : h->highest;
    if (standalone) gf(max_lvl);
    for (int lvl = max_lvl; lvl >= min_lvl + 1; --lvl) { smooth_boxes(lvl, h->params.n_cycle_down); update_coarse(lvl); }
    const int sub = (h->params.smoother_type == 2) ? 2 : 1;
    if (h->L(min_lvl).nb > 0 || !multi())
      pend.push_back({TOP_COARSE_SOLVE, li(min_lvl), (h->params.n_cycle_up + h->params.n_cycle_down) * sub,
                      h->params.max_coarse_cycles});
    for (int lvl = min_lvl + 1; lvl <= max_lvl; ++lvl) {
      correct_children_onto(lvl, h->params.n_cycle_up); smooth_boxes(lvl, h->params.n_cycle_up);
    }
    if (want_res) {
      step(S_CLEAR_SLOT, 0, SCAL_MAXRES);
      for (int lvl = min_lvl; lvl <= max_lvl; ++lvl) emit(lvl, TOP_RESID, S_RESID, SCAL_MAXRES);
      if (multi()) step(S_ALLREDUCE_MAX, 0, SCAL_MAXRES);
      if (res_slot != SCAL_MAXRES) step(S_COPY_SLOT, 0, res_slot, SCAL_MAXRES);
    }
    if (h->params.subtract_mean) subtract_mean(IPHI, 1);
  }
  // mg_fas_fmg :84-147
  void fmg(bool have_guess, bool want_res) {
    if (!have_guess) {
      for (int lvl = h->highest; lvl >= h->lowest; --lvl) { step(S_ZERO_PHI, lvl); h->gz[lvl - OMG_LVL_LO] = true; }
    }
    gf(h->highest);
    for (int lvl = h->highest; lvl >= h->lowest + 1; --lvl) update_coarse(lvl);
    if (h->params.subtract_mean) subtract_mean(IRHS, 0);
    for (int lvl = h->lowest; lvl <= h->highest; ++lvl) {
      emit(lvl, TOP_COPY_OLD, S_COPY_OLD);
      if (!h->L(lvl).h_oldpl.empty()) emit(lvl, TOP_OLDPL, S_OLDPL);
      if (lvl > h->lowest) correct_children_onto(lvl, h->params.n_cycle_down);     // the V-cycle from lvl starts with its down-smoothing
      vcycle(true, lvl, false, want_res && lvl == h->highest, SCAL_MAXRES);
    }
  }
};

// A bounded in-kernel wait for a peer's halo signal expired (omg_comm.cu wait_flag): the cycle went on with stale
// planes, so everything computed since is void.  The error word lives in mapped host memory: checking it is a host
// load.  Called at every host synchronisation point and at the start of every entry point; once raised the handle
// stays poisoned (the reference's MPI_Waitall would have blocked for ever or aborted the job).
void check_p2p(omg_handle* h) {
  if (!h->comm) return;
  const int code = omg_comm_p2p_error(h->comm);
  if (!code) return;
  const int ch = omg_p2p_err_channel(code);
  char buf[256];
  std::snprintf(buf, sizeof buf, "halo exchange timed out: rank %d waited more than %.0f s for rank %d on %s (peer gone or too far behind; "
                "OMG_P2P_TIMEOUT_S sets the bound, 0 = none); results of this handle are invalid from here on",
                h->my_rank, omg_comm_p2p_timeout_ns(h->comm) * 1e-9, omg_p2p_err_peer(code),
                ch == OMG_P2P_CHANNELS - 1 ? "the teardown rendezvous" : ("level " + std::to_string(ch + OMG_LVL_LO)).c_str());
  fail(buf);
}

// ---- executor ------------------------------------------------------------------------------------
void need_comm(omg_handle* h) {
  if (!h->comm) fail("several ranks but no communicator: call omg_comm_init (mg_comm_init) after omg_set_tree");
  if (omg_comm_size(h->comm) != h->n_cpu || omg_comm_rank(h->comm) != h->my_rank) fail("communicator rank/size differ from mg%my_rank / mg%n_cpu");
}

void run_exchange(omg_handle* h, const PhasePlan& X, const double* send, double* recv, size_t item, cudaStream_t s) {
  need_comm(h);
  std::vector<OmgXfer> segs;
  for (const PhasePlan::Seg& g : X.seg)
    segs.push_back({g.peer, send + (size_t)g.send_off * item, (size_t)g.send_n * item, recv + (size_t)g.recv_off * item, (size_t)g.recv_n * item});
  try { omg_comm_exchange(h->comm, segs.data(), (int)segs.size(), s); }
  catch (const std::exception& e) { fail(e.what()); }
  h->exchanges++;
}

// mg_fill_ghost_cells_lvl, communication part (src/m_ghost_cells.f90:142-168): faces owned by other ranks
// arrive directly in this level's plane pool
// Two transports.  put (peer-mapped pools): the pack kernel stores the faces straight into the
// receivers' plane pools over NVLink, then one tiny kernel signals the peers and waits for their signal.
// send/recv: pack into a send buffer, one ncclSend/ncclRecv pair per peer.
bool ghost_by_put(omg_handle* h, int lvl) { return h->p2p && h->L(lvl).xg.shared; }
void pack_ghost(omg_handle* h, int lvl, int gz, cudaStream_t s) {
  LevelHost& L = h->L(lvl);
  const LevelDev& Lc = lvl > h->lowest ? h->L(lvl - 1).dev : L.dev;
  if (!L.xg.n_send) return;
  if (ghost_by_put(h, lvl))
    omgk::put_ghost(L.dev, Lc, h->pdev, L.xg.d_sdesc.p, L.xg.d_dst.p, L.xg.n_send, L.xg.d_peer_base.p, L.xg.d_peer_stride.p,
                    h->d_epoch.p + (lvl - OMG_LVL_LO), gz, s), h->puts++;
  else omgk::pack_ghost(L.dev, Lc, h->pdev, L.xg.d_sdesc.p, L.xg.n_send, L.xg.d_send.p, gz, s);
  h->launches++;
}
void transfer_ghost(omg_handle* h, int lvl, cudaStream_t s) {
  LevelHost& L = h->L(lvl);
  if (ghost_by_put(h, lvl)) {
    need_comm(h);
    try { omg_comm_signal_wait(h->comm, lvl - OMG_LVL_LO, L.xg.peers.data(), (int)L.xg.peers.size(), h->d_epoch.p + (lvl - OMG_LVL_LO), 1, s); }
    catch (const std::exception& e) { fail(e.what()); }
    h->exchanges++; h->launches++;
  } else run_exchange(h, L.xg, L.xg.d_send.p, L.d_gpl.p, (size_t)L.nc * L.nc, s);
}
void exchange_ghost(omg_handle* h, int lvl, int gz, cudaStream_t s) {
  pack_ghost(h, lvl, gz, s);
  transfer_ghost(h, lvl, s);
}
// mg_restrict_lvl, communication part (src/m_restrict.f90:94-108); both variables of update_coarse in one message
void exchange_restrict(omg_handle* h, int lvl, int iv0, int iv1, cudaStream_t s) {
  LevelHost& L = h->L(lvl);
  const int hn = L.nc / 2;
  const size_t item = (size_t)(iv1 >= 0 ? 2 : 1) * hn * hn * hn;
  if (L.xr.n_send) { omgk::pack_restrict(L.dev, L.xr.d_sdesc.p, L.xr.n_send, L.xr.d_send.p, iv0, iv1, s); h->launches++; }
  run_exchange(h, L.xr, L.xr.d_send.p, L.xr.d_recv.p, item, s);
  if (L.xr.n_recv) { omgk::unpack_restrict(h->L(lvl - 1).dev, L.xr.d_rdesc.p, L.xr.n_recv, L.xr.d_recv.p, hn, iv0, iv1, s); h->launches++; }
}
// mg_prolong, communication part (src/m_prolong.f90:65-78): coarse correction patches instead of fine blocks
void exchange_prolong(omg_handle* h, int lvl, cudaStream_t s) {
  LevelHost& L = h->L(lvl);
  const int pw = L.nc / 2 + 2;
  if (L.xp.n_send) { omgk::pack_prolong(h->L(lvl - 1).dev, h->pdev, L.xp.d_sdesc.p, L.xp.n_send, L.xp.d_send.p, L.nc, s); h->launches++; }
  run_exchange(h, L.xp, L.xp.d_send.p, L.xp.d_recv.p, (size_t)pw * pw * pw, s);
}

void run_submean(omg_handle* h, int iv, int with_ghosts) {
  // subtract_mean + get_sum, src/m_multigrid.f90:245-294
  cudaStream_t s = h->stream;
  double* acc = h->d_scal.p + SCAL_MEAN;
  omgk::fill(acc, 1, 0.0, s); h->launches++;
  for (int lvl = std::max(1, h->lowest); lvl <= h->highest; ++lvl) {
    LevelHost& L = h->L(lvl);
    if (L.h_leaves.empty()) continue;
    const double w = L.dev.dr[0] * L.dev.dr[1] * L.dev.dr[2];
    omgk::box_sums(L.dev, iv, w, h->d_partial.p, s);
    omgk::final_sum(h->d_partial.p, (int)L.h_leaves.size(), acc, s);
    h->launches += 2;
  }
  if (h->n_cpu > 1) { need_comm(h); omg_comm_allreduce(h->comm, acc, OMG_RED_SUM, s); }
  if (1 >= h->lowest && 1 <= h->highest) {
    LevelHost& L1 = h->L(1);
    const double nc = (double)(h->box_size * h->box_size * h->box_size);
    const double volume = nc * (L1.dev.dr[0] * L1.dev.dr[1] * L1.dev.dr[2]) * (double)h->lvl_ids[1 - h->lowest].size();
    omgk::div_scalar(acc, volume, s); h->launches++;
  }
  for (LevelHost& L : h->levels) {
    omgk::sub_scalar(L.d_var[iv].p, L.d_var[iv].n, acc, s); h->launches++;
    if (with_ghosts && iv == IPHI && L.d_gpl.n) { omgk::sub_scalar(L.d_gpl.p, L.d_gpl.n, acc, s); h->launches++; }
  }
}

void exec_steps(omg_handle* h, Program& P) {
  cudaStream_t s = h->stream;
  const ParamsDev& PD = h->pdev;
  omgf::set_pdl(h->pdl && h->n_cpu == 1 && !h->profile);
  omgf::set_ring(h->ring);
  omgf::set_store_res(h->store_fine_residual);
  auto cells_of = [](const LevelHost& L) { return (int64_t)L.nb * L.nc * L.nc * L.nc; };
  auto fast = [&](const LevelHost& L, int n = 1) { h->fast_launches += n; h->fast_cells += n * cells_of(L); };
  auto general = [&](const LevelHost& L, int n = 1) { h->general_launches += n; h->general_cells += n * cells_of(L); };
  for (const Step& st : P.steps) {
    const bool prof = h->profile && st.kind != S_SUBMEAN && st.kind != S_CLEAR_SLOT && st.kind != S_COPY_SLOT &&
                      st.kind != S_ZERO_PHI && st.kind != S_HALO_FORK && st.kind != S_HALO_JOIN;
    cudaEvent_t pe0 = nullptr, pe1 = nullptr;
    if (prof) { CK(cudaEventCreate(&pe0)); CK(cudaEventCreate(&pe1)); CK(cudaEventRecord(pe0, s)); }
    cudaEvent_t te0 = nullptr, te1 = nullptr;
    if (h->trace) { CK(cudaEventCreate(&te0)); CK(cudaEventCreate(&te1)); CK(cudaEventRecord(te0, s)); }
    switch (st.kind) {
      case S_SMOOTH: {
        LevelHost& L = h->L(st.l);
        if (st.b == 3) {   // in-kernel exchange: whole level in one launch
          if (L.nb) {
            // In-kernel exchange: the one-tile-per-CTA kernel, not the persistent ring (option "ring_fused" = 1 switches it on).
            // A ring CTA that reaches a partition-face tile stalls its whole pipeline for the wait and the remote stores;
            // measured at 2 GPUs, 512^3 each (profiles/r02p_*): ring with partition-face boxes first 4.899 ms per cycle,
            // last 4.999, natural order 5.943 -- one tile per CTA 4.793, although its half-sweep alone is 0.279 vs 0.268 ms.
            // Launch order: partition-face boxes first on small levels (their faces leave early), natural order on big ones.
            const bool ring = h->ring_fused && omgf::smooth_uses_ring(L.dev, PD, L.nb);
            omgf::set_ring(ring ? h->ring : 0);
            const int* order = ring ? (h->ring_order == 1 ? L.d_launch_ring.p : h->ring_order == 2 ? nullptr : L.d_launch.p)
                                    : (L.nb <= h->bnd_first_boxes ? L.d_launch.p : nullptr);
            const bool ok_fused = omgf::smooth(L.dev, PD, st.a, L.all_fast, order, L.nb, h->put_fence == 2 ? 2 : 1, s);
            omgf::set_ring(h->ring);
            if (!ok_fused) fail("fused smoother does not apply to this level");
            h->launches++; h->exchanges++; h->fused_smooths++; fast(L);
          }
          // emulated ranks share one GPU: a kernel must never wait for a kernel that may not be running, so the ranks
          // meet on the host and order their streams with events; the in-kernel waits then return at once
          if (!omg_comm_graph_safe(h->comm)) {
            try { omg_comm_signal_wait(h->comm, st.l - OMG_LVL_LO, L.xg.peers.data(), (int)L.xg.peers.size(), h->d_epoch.p + (st.l - OMG_LVL_LO), 0, s); }
            catch (const std::exception& e) { fail(e.what()); }
          }
          break;
        }
        const int* list = st.b == 1 ? L.d_bnd.p : st.b == 2 ? L.d_int.p : nullptr;
        const int nlist = st.b == 1 ? (int)L.h_bnd.size() : st.b == 2 ? (int)L.h_int.size() : 0;
        if (st.b && nlist == 0) break;
        if (PD.smoother == 1) {   // mg_smoother_gs: ghost snapshot, then the in-box sweep (shared-memory staged where it applies)
          omgk::gs_snapshot(L.dev, PD, s); h->launches++;
          if (h->fast_kernels && L.nb && omgf::gs_sweep(L.dev, PD, s)) fast(L);
          else { omgk::gs_sweep(L.dev, PD, s); general(L); }
        }
        else if (h->fast_kernels && omgf::smooth(L.dev, PD, st.a, L.all_fast, st.b ? list : nullptr, nlist, 0, s)) fast(L);
        else { omgk::smooth(L.dev, PD, st.a, st.b ? list : nullptr, nlist, s); general(L); }
        h->launches++;
      } break;
      case S_RESID: {
        LevelHost& L = h->L(st.l);
        double* slot = st.a >= 0 ? h->d_scal.p + st.a : nullptr;
        if (h->fast_kernels && omgf::resid_max(L.dev, PD, L.all_fast, slot, s)) fast(L);
        else { omgk::resid(L.dev, PD, slot, s); general(L); }
        h->launches++;
      } break;
      case S_RESTRICT: omgk::restrict2(h->L(st.l).dev, h->L(st.l - 1).dev, st.a, st.b, s); h->launches++; general(h->L(st.l)); break;
      case S_RESID_RESTRICT: {
        LevelHost& L = h->L(st.l);
        if (h->fast_kernels && omgf::resid_restrict(L.dev, h->L(st.l - 1).dev, PD, L.all_fast, s)) fast(L);
        else {
          omgk::resid(L.dev, PD, nullptr, s); omgk::restrict2(L.dev, h->L(st.l - 1).dev, IPHI, IRES, s); h->launches++;
          general(L, 2);
        }
        h->launches++;
      } break;
      case S_CRHS: {
        LevelHost& L = h->L(st.l);
        if (h->fast_kernels && omgf::crhs(L.dev, PD, L.all_fast, s)) fast(L);
        else { omgk::crhs(L.dev, PD, s); general(L); }
        h->launches++;
      } break;
      case S_OLDPL: omgk::oldpl(h->L(st.l).dev, PD, s); h->launches++; break;
      case S_CORR_PROLONG: {
        LevelHost& L = h->L(st.l);
        if (st.a == 3) {   // in-kernel exchange of the corrected faces (see Builder::correct_children_onto)
          if (L.nb) {
            if (!omgf::corr_prolong(L.dev, h->L(st.l - 1).dev, PD, h->L(st.l - 1).all_fast, h->put_fence == 2 ? 2 : 1, s)) fail("fused prolongation does not apply to this level");
            h->launches++; h->exchanges++; h->fused_prolongs++; fast(L);
          }
          if (!omg_comm_graph_safe(h->comm)) {      // emulated ranks: host rendezvous instead of in-kernel waits (see S_SMOOTH)
            try { omg_comm_signal_wait(h->comm, st.l - OMG_LVL_LO, L.xg.peers.data(), (int)L.xg.peers.size(), h->d_epoch.p + (st.l - OMG_LVL_LO), 0, s); }
            catch (const std::exception& e) { fail(e.what()); }
          }
          break;
        }
        if (h->fast_kernels && omgf::corr_prolong(L.dev, h->L(st.l - 1).dev, PD, h->L(st.l - 1).all_fast, 0, s)) fast(L);
        else { omgk::corr_prolong(L.dev, h->L(st.l - 1).dev, PD, s); general(L); }
        h->launches++;
      } break;
      case S_RB_FILL: omgk::rb_fill(h->L(st.l).dev, h->L(st.l - 1).dev, PD, 0, st.a, s); h->launches++; break;
      case S_COPY_OLD: { LevelHost& L = h->L(st.l); omgk::copy(L.d_var[IOLD].p, L.d_var[IPHI].p, L.d_var[IPHI].n, s); h->launches++; } break;
      case S_TAIL: {
        // stage the tail levels in shared memory when all four variables of all of them fit (option "tail_smem")
        int n_stage = 0; size_t bytes = 0;
        if (h->tail_smem) for (LevelHost& L : h->levels) { if (!L.tail) break; ++n_stage; bytes += (size_t)4 * L.nb * L.dev.bs * sizeof(double) + (size_t)32 * L.nb * sizeof(int); }   // variables + face / box tables (k_tail)
        // levels beyond the single-CTA limit in the tail group: only the cluster interpreter can take them
        bool big = false;
        for (LevelHost& L : h->levels) { if (!L.tail) break; if ((long)L.nb * L.nc * L.nc * L.nc > h->tail_cells) big = true; }
        if (big || (h->tail_cluster >= 2 && h->tail_cluster_always && cluster_tail_ok(h))) {
          if (!cluster_tail_ok(h) || !omgk::tail_cluster(h->d_levels.p, PD, P.d_ops.p + st.a, st.b, h->d_scal.p, h->tail_cluster, s))
            fail("cluster tail: launch failed (option tail_cluster = 0 switches it off)");
        } else omgk::tail(h->d_levels.p, PD, P.d_ops.p + st.a, st.b, h->d_scal.p, n_stage, bytes, s);
        h->launches++;
      } break;
      case S_ZERO_PHI: {
        LevelHost& L = h->L(st.l);
        if (L.d_var[IPHI].n) CK(cudaMemsetAsync(L.d_var[IPHI].p, 0, L.d_var[IPHI].n * sizeof(double), s));
        if (L.d_gpl.n) CK(cudaMemsetAsync(L.d_gpl.p, 0, L.d_gpl.n * sizeof(double), s));
      } break;
      case S_SUBMEAN: run_submean(h, st.a, st.b); break;
      case S_CLEAR_SLOT: CK(cudaMemsetAsync(h->d_scal.p + st.a, 0, sizeof(double), s)); break;
      case S_COPY_SLOT: CK(cudaMemcpyAsync(h->d_scal.p + st.a, h->d_scal.p + st.b, sizeof(double), cudaMemcpyDeviceToDevice, s)); break;
      case S_HALO: exchange_ghost(h, st.l, st.a, s); break;
      case S_HALO_FORK: {
        // pack on the compute stream, then fork: the NCCL kernel needs (nearly) empty SMs, so it must be dispatched
        // before the interior half-sweep floods the machine -- the exchange stream has the higher priority and both
        // kernels become ready at the same event
        pack_ghost(h, st.l, st.a, s);
        CK(cudaEventRecord(h->ev_fork, s));
        CK(cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
        transfer_ghost(h, st.l, h->stream2);
        CK(cudaEventRecord(h->ev_join, h->stream2));
      } break;
      case S_HALO_JOIN: CK(cudaStreamWaitEvent(s, h->ev_join, 0)); break;
      case S_PUT_WAIT: {
        LevelHost& L = h->L(st.l);
        if (omg_comm_graph_safe(h->comm)) {    // NCCL back end: one single-CTA kernel polling the peers' flags
          try { omg_comm_signal_wait(h->comm, st.l - OMG_LVL_LO, L.xg.peers.data(), (int)L.xg.peers.size(), h->d_epoch.p + (st.l - OMG_LVL_LO), 0, s); }
          catch (const std::exception& e) { fail(e.what()); }
          h->launches++;
        }
      } break;
      case S_XRESTRICT: exchange_restrict(h, st.l, st.a, st.b, s); break;
      case S_XPROLONG: exchange_prolong(h, st.l, s); break;
      case S_ALLREDUCE_MAX: need_comm(h); omg_comm_allreduce(h->comm, h->d_scal.p + st.a, OMG_RED_MAX, s); break;
      default: fail("exec_steps: unknown step");
    }
    if (prof) { CK(cudaEventRecord(pe1, s)); h->prof_events.push_back({st.kind * 2 + (st.kind != S_TAIL && st.l == h->highest ? 1 : 0), {pe0, pe1}}); }
    if (h->trace) { CK(cudaEventRecord(te1, s)); h->trace_events.push_back({st.kind, st.l, st.a, st.b, te0, te1}); }
  }
  CK(cudaGetLastError());
}

void run_program(omg_handle* h, Program& P) {
  if (P.d_ops.n != P.ops.size() || (!P.d_ops.p && !P.ops.empty())) {
    P.d_ops.upload(P.ops, h->stream);
    CK(cudaStreamSynchronize(h->stream));
  }
  const bool graph_ok = h->use_graph && !h->profile && !h->trace && (h->n_cpu == 1 || (h->comm && h->graph_multi && omg_comm_graph_safe(h->comm)));
  if (!graph_ok) { exec_steps(h, P); return; }
  if (!P.graph) {
    cudaGraph_t g = nullptr;
    const int64_t l0 = h->launches, f0 = h->fast_launches, g0 = h->general_launches, fc0 = h->fast_cells, gc0 = h->general_cells;
    CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    try { exec_steps(h, P); } catch (...) { cudaStreamEndCapture(h->stream, &g); if (g) cudaGraphDestroy(g); throw; }
    CK(cudaStreamEndCapture(h->stream, &g));
    CK(cudaGraphInstantiate(&P.graph, g, 0));
    CK(cudaGraphDestroy(g));
    P.launches_per_run = h->launches - l0; P.fast_per_run = h->fast_launches - f0; P.general_per_run = h->general_launches - g0;
    P.fastc_per_run = h->fast_cells - fc0; P.generalc_per_run = h->general_cells - gc0;
    h->launches = l0; h->fast_launches = f0; h->general_launches = g0; h->fast_cells = fc0; h->general_cells = gc0;
  }
  CK(cudaGraphLaunch(P.graph, h->stream));
  h->launches += P.launches_per_run; h->fast_launches += P.fast_per_run; h->general_launches += P.general_per_run;
  h->fast_cells += P.fastc_per_run; h->general_cells += P.generalc_per_run;
  h->graph_launches++;
}

// totals "mg total V-cycle" / "mg total FMG cycle": CUDA events around the program, read lazily
void drain_totals(omg_handle* h, bool wait) {
  size_t keep = 0;
  for (auto& e : h->total_events) {
    if (!wait && cudaEventQuery(e.second.second) != cudaSuccess) { h->total_events[keep++] = e; continue; }
    if (wait) CK(cudaEventSynchronize(e.second.second));
    float ms = 0;
    if (cudaEventElapsedTime(&ms, e.second.first, e.second.second) == cudaSuccess) h->timer_acc[e.first] += ms * 1e-3;
    cudaEventDestroy(e.second.first); cudaEventDestroy(e.second.second);
  }
  h->total_events.resize(keep);
}
void timed_run(omg_handle* h, Program& P, int timer) {
  drain_totals(h, h->total_events.size() > 64);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaEventRecord(e0, h->stream));
  try { run_program(h, P); } catch (...) { cudaEventDestroy(e0); cudaEventDestroy(e1); throw; }
  CK(cudaEventRecord(e1, h->stream));
  h->total_events.push_back({timer, {e0, e1}});
}

Program& get_program(omg_handle* h, const std::string& key, bool& fresh) {
  auto it = h->programs.find(key);
  fresh = (it == h->programs.end());
  return h->programs[key];
}

void check_supported(omg_handle* h) {
  const omg_params& p = h->params;
  if (p.operator_type < 1 || p.operator_type > 5) fail("mg_set_methods: unknown operator");
  if (p.geometry_type != 1) fail("set_methods: unsupported geometry");
  if (p.smoother_type != 1 && p.smoother_type != 2) fail("set_methods: unsupported smoother type");
  if ((p.operator_type == OP_VLPL || p.operator_type == OP_VHELMH) && h->n_extra < 1) fail("set_methods: mg%n_extra_vars == 0");
  if (p.operator_type == OP_AHELMH && h->n_extra < 3) fail("ahelmholtz_set_methods: mg%n_extra_vars < 3");
}

void fas_vcycle(omg_handle* h, int has_highest, int highest_lvl, int standalone, int want_res, double* max_res) {
  CK(cudaSetDevice(h->device));
  check_p2p(h);
  sync_meta(h);
  check_supported(h);
  if (has_highest && (highest_lvl < h->lowest || highest_lvl > h->highest)) fail("mg_fas_vcycle: highest_lvl out of range");
  char key[64];
  std::snprintf(key, sizeof key, "v:%d:%d:%d:%d", has_highest, has_highest ? highest_lvl : 0, standalone, want_res);
  bool fresh;
  Program& P = get_program(h, key, fresh);
  if (fresh) {
    for (int l = 0; l < OMG_NLVL; ++l) h->gz[l] = false;
    Builder B{h, P, {}};
    B.vcycle(has_highest != 0, highest_lvl, standalone != 0, want_res != 0, SCAL_MAXRES);
    B.flush();
  }
  timed_run(h, P, 0);
  if (want_res && max_res) {
    CK(cudaMemcpyAsync(max_res, h->d_scal.p + SCAL_MAXRES, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    check_p2p(h);
  }
}

void fas_fmg(omg_handle* h, int have_guess, int want_res, double* max_res) {
  CK(cudaSetDevice(h->device));
  check_p2p(h);
  sync_meta(h);
  check_supported(h);
  char key[64];
  std::snprintf(key, sizeof key, "f:%d:%d", have_guess, want_res);
  bool fresh;
  Program& P = get_program(h, key, fresh);
  if (fresh) {
    for (int l = 0; l < OMG_NLVL; ++l) h->gz[l] = false;
    Builder B{h, P, {}};
    B.fmg(have_guess != 0, want_res != 0);
    B.flush();
  }
  timed_run(h, P, 1);
  if (want_res && max_res) {
    CK(cudaMemcpyAsync(max_res, h->d_scal.p + SCAL_MAXRES, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    check_p2p(h);
  }
}

// ---- staging for push / pull -----------------------------------------------------------------------
void ensure_stage(omg_handle* h, size_t slab) {
  const size_t want_boxes = std::max<size_t>(1, std::min<size_t>(4096, ((size_t)48 << 20) / (slab * sizeof(double))));
  const size_t need = want_boxes * slab;
  if (h->stage_doubles < need) {
    for (int q = 0; q < 2; ++q) {
      if (h->h_stage[q]) cudaFreeHost(h->h_stage[q]);
      if (h->d_stage[q]) cudaFree(h->d_stage[q]);
      CK(cudaMallocHost(&h->h_stage[q], need * sizeof(double)));
      CK(cudaMalloc(&h->d_stage[q], need * sizeof(double)));
      if (!h->d_boxidx[q]) CK(cudaMalloc(&h->d_boxidx[q], 4096 * sizeof(int)));
      if (!h->h_boxidx[q]) CK(cudaMallocHost(&h->h_boxidx[q], 4096 * sizeof(int)));
      if (!h->stage_ev[q]) CK(cudaEventCreateWithFlags(&h->stage_ev[q], cudaEventDisableTiming));
    }
    h->stage_doubles = need;
  }
  h->stage_boxes = (int)std::min<size_t>(4096, h->stage_doubles / slab);
}

void collect_profile(omg_handle* h) {
  if (h->prof_events.empty()) return;
  CK(cudaStreamSynchronize(h->stream));
  for (auto& e : h->prof_events) {
    float ms = 0; CK(cudaEventElapsedTime(&ms, e.second.first, e.second.second));
    const int kind = e.first >> 1;
    if (e.first & 1) { auto& acc = h->prof_acc[kind]; acc.first += ms; acc.second += 1; }
    // phase timers of the reference: smoother, smoother g.c., coarse, correct, update coarse
    int t = -1;
    switch (kind) {
      case S_SMOOTH: t = 2; break;
      case S_HALO: case S_RB_FILL: t = 3; break;
      case S_TAIL: case S_COARSE_SOLVE: t = 4; break;
      case S_CORR_PROLONG: case S_XPROLONG: t = 5; break;
      case S_RESID_RESTRICT: case S_RESID: case S_RESTRICT: case S_XRESTRICT: case S_CRHS: case S_OLDPL: case S_COPY_OLD: t = 6; break;
      default: break;
    }
    if (t >= 0) h->timer_acc[t] += ms * 1e-3;
    cudaEventDestroy(e.second.first); cudaEventDestroy(e.second.second);
  }
  h->prof_events.clear();
}

int ghost_mode_for(int v) { return v == IPHI ? (int)GM_PHI : v == IOLD ? (int)GM_OLD : v >= 4 ? (int)GM_STORED : -1; }

// copy interior (+ six faces) between a slab and a user cc(:,:,:,iv) array
void scatter_slab(double* dst, const double* src, int nc, bool faces) {
  const int n2 = nc + 2;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j)
    std::memcpy(dst + ((size_t)k * n2 + j) * n2 + 1, src + ((size_t)k * n2 + j) * n2 + 1, sizeof(double) * nc);
  if (!faces) return;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) {
    const size_t r = ((size_t)k * n2 + j) * n2;
    dst[r] = src[r]; dst[r + nc + 1] = src[r + nc + 1];
  }
  for (int k = 1; k <= nc; ++k) {
    std::memcpy(dst + ((size_t)k * n2 + 0) * n2 + 1, src + ((size_t)k * n2 + 0) * n2 + 1, sizeof(double) * nc);
    std::memcpy(dst + ((size_t)k * n2 + nc + 1) * n2 + 1, src + ((size_t)k * n2 + nc + 1) * n2 + 1, sizeof(double) * nc);
  }
  for (int j = 1; j <= nc; ++j) {
    std::memcpy(dst + ((size_t)0 * n2 + j) * n2 + 1, src + ((size_t)0 * n2 + j) * n2 + 1, sizeof(double) * nc);
    std::memcpy(dst + ((size_t)(nc + 1) * n2 + j) * n2 + 1, src + ((size_t)(nc + 1) * n2 + j) * n2 + 1, sizeof(double) * nc);
  }
}

// transfer a list of (local box, host pointer) pairs of one level; packed: host slabs are contiguous from base
void push_level(omg_handle* h, LevelHost& L, int v, const std::vector<int>& boxes, const double* const* ptrs, const double* packed, int shell = 1) {
  const size_t slab = (size_t)(L.nc + 2 * shell) * (L.nc + 2 * shell) * (L.nc + 2 * shell);
  ensure_stage(h, slab);
  const int chunk = h->stage_boxes;
  int q = 0;
  for (size_t c0 = 0; c0 < boxes.size(); c0 += chunk, q ^= 1) {
    const int n = (int)std::min<size_t>(chunk, boxes.size() - c0);
    CK(cudaEventSynchronize(h->stage_ev[q]));
    const double* src;
    if (packed) src = packed + c0 * slab;
    else {
      for (int m = 0; m < n; ++m) std::memcpy(h->h_stage[q] + (size_t)m * slab, ptrs[c0 + m], slab * sizeof(double));
      src = h->h_stage[q];
    }
    std::memcpy(h->h_boxidx[q], boxes.data() + c0, n * sizeof(int));
    CK(cudaMemcpyAsync(h->d_stage[q], src, (size_t)n * slab * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->d_boxidx[q], h->h_boxidx[q], n * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    omgk::unpack(L.dev, v, h->d_boxidx[q], n, h->d_stage[q], v >= 4, shell, h->stream);
    CK(cudaEventRecord(h->stage_ev[q], h->stream));
    h->launches++; h->bytes_h2d += (int64_t)n * slab * sizeof(double);
  }
  CK(cudaStreamSynchronize(h->stream));
  check_p2p(h);
}

void pull_level(omg_handle* h, LevelHost& L, int v, const std::vector<int>& boxes, double* const* ptrs, double* packed, int with_ghosts, int shell = 1) {
  const size_t slab = (size_t)(L.nc + 2 * shell) * (L.nc + 2 * shell) * (L.nc + 2 * shell);
  ensure_stage(h, slab);
  const int chunk = h->stage_boxes;
  const int gm = with_ghosts ? ghost_mode_for(v) : -1;
  struct Pending { size_t c0; int n, q; };
  Pending prev{0, 0, -1};
  auto drain = [&](const Pending& p) {
    if (p.q < 0) return;
    CK(cudaEventSynchronize(h->stage_ev[p.q]));
    if (!packed) for (int m = 0; m < p.n; ++m) scatter_slab(ptrs[p.c0 + m], h->h_stage[p.q] + (size_t)m * slab, L.nc, gm >= 0);
  };
  int q = 0;
  for (size_t c0 = 0; c0 < boxes.size(); c0 += chunk, q ^= 1) {
    const int n = (int)std::min<size_t>(chunk, boxes.size() - c0);
    std::memcpy(h->h_boxidx[q], boxes.data() + c0, n * sizeof(int));
    CK(cudaMemcpyAsync(h->d_boxidx[q], h->h_boxidx[q], n * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    omgk::pack(L.dev, h->pdev, v, h->d_boxidx[q], n, h->d_stage[q], gm, shell, h->stream);
    double* dst = packed ? packed + c0 * slab : h->h_stage[q];
    CK(cudaMemcpyAsync(dst, h->d_stage[q], (size_t)n * slab * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaEventRecord(h->stage_ev[q], h->stream));
    h->launches++; h->bytes_d2h += (int64_t)n * slab * sizeof(double);
    drain(prev);
    prev = {c0, n, q};
  }
  drain(prev);
  CK(cudaStreamSynchronize(h->stream));
  check_p2p(h);
}

int check_iv(omg_handle* h, int iv) {
  if (iv < 1 || iv > h->nvar) fail("variable index out of range");
  return iv - 1;
}

void push_pull_ids(omg_handle* h, int iv, int n, const int32_t* ids, const double* const* cptrs, double* const* ptrs, int with_ghosts, bool push) {
  CK(cudaSetDevice(h->device));
  sync_meta(h);
  const int v = check_iv(h, iv);
  // group by level, preserving order
  std::map<int, std::pair<std::vector<int>, std::vector<size_t>>> by_lvl;
  for (int m = 0; m < n; ++m) {
    const int id = ids[m];
    if (id < 1 || id > h->n_boxes || h->loc[id] < 0) fail("push/pull: box not owned by this rank");
    auto& e = by_lvl[h->blvl[id - 1]];
    e.first.push_back(h->loc[id]); e.second.push_back(m);
  }
  for (auto& kv : by_lvl) {
    LevelHost& L = h->L(kv.first);
    if (push) {
      std::vector<const double*> p; for (size_t m : kv.second.second) p.push_back(cptrs[m]);
      push_level(h, L, v, kv.second.first, p.data(), nullptr);
    } else {
      std::vector<double*> p; for (size_t m : kv.second.second) p.push_back(ptrs[m]);
      pull_level(h, L, v, kv.second.first, p.data(), nullptr, with_ghosts);
    }
  }
}

// boxes of a level taking part in a packed transfer: all local boxes, or the local leaves only (in box order)
std::vector<int> packed_boxes(const LevelHost& L, bool leaves_only) {
  if (leaves_only) return L.h_leaves;
  std::vector<int> boxes(L.nb);
  for (int b = 0; b < L.nb; ++b) boxes[b] = b;
  return boxes;
}

// shell = 1: slabs in the reference's cc layout (ghost shell included); shell = 0: interior cells only
void push_pull_packed(omg_handle* h, int iv, int lvl_lo, int lvl_hi, const double* cbuf, double* buf, int64_t n_doubles, int with_ghosts, bool push,
                      int shell = 1, bool leaves_only = false) {
  CK(cudaSetDevice(h->device));
  sync_meta(h);
  const int v = check_iv(h, iv);
  lvl_lo = std::max(lvl_lo, h->lowest); lvl_hi = std::min(lvl_hi, h->highest);
  if (lvl_lo > lvl_hi) fail("packed transfer: empty level range");
  int64_t off = 0;
  for (int lvl = lvl_lo; lvl <= lvl_hi; ++lvl) {
    LevelHost& L = h->L(lvl);
    const int64_t edge = L.nc + 2 * shell, slab = edge * edge * edge;
    const std::vector<int> boxes = packed_boxes(L, leaves_only);
    if (boxes.empty()) continue;
    if (off + slab * (int64_t)boxes.size() > n_doubles) fail("packed transfer: buffer too small");
    if (push) push_level(h, L, v, boxes, nullptr, cbuf + off, shell);
    else pull_level(h, L, v, boxes, nullptr, buf + off, with_ghosts, shell);
    off += slab * (int64_t)boxes.size();
  }
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
#define OMG_TRY(h, ...)                                                    \
  try { __VA_ARGS__; return 0; }                                           \
  catch (const OmgError& e) { if (h) (h)->err = e.msg; return 1; }         \
  catch (const std::exception& e) { if (h) (h)->err = e.what(); return 2; }

extern "C" {

const char* omg_version(void) { return "octree-mg_b200 0.1 (sm_100a)"; }

int omg_create(int device, omg_handle** out) {
  if (!out) return 1;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) return 3;   // no CPU fallback
  omg_handle* h = new omg_handle();
  h->device = device;
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return 3; }
  cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
  {
    // the exchange stream outranks the compute stream: its pack / NCCL kernels must be dispatched as soon as SM
    // resources free up, not after the interior half-sweep has drained
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    cudaStreamCreateWithPriority(&h->stream2, cudaStreamNonBlocking, prio_hi);
  }
  cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming); cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
  for (int v = 0; v < OMG_MAXV; ++v) for (int f = 0; f < 6; ++f) { h->bc_type[v][f] = -10; h->bc_value[v][f] = 0.0; }
  // defaults of type(mg_t), src/m_data_structures.f90:300-327
  h->params.operator_type = 1; h->params.smoother_type = 1; h->params.geometry_type = 1;
  h->params.n_cycle_down = 2; h->params.n_cycle_up = 2; h->params.max_coarse_cycles = 1000;
  h->params.residual_coarse_abs = 1e-8; h->params.residual_coarse_rel = 1e-8;
  h->have_params = true;
  // tuning / bisecting aid: OMG_RING_DEFAULT overrides the default of option "ring" (omg_set_option still wins)
  if (const char* e = getenv("OMG_RING_DEFAULT")) { if (*e) h->ring = atoi(e); }
  *out = h;
  return 0;
}

int omg_destroy(omg_handle* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  // graph executables first: NCCL keeps a communicator alive (ncclCommDestroy does not return) while a captured graph
  // that contains its send / recv / all-reduce nodes still exists -- which is what blocked every multi-rank teardown
  // (gpurun_out r02n: both ranks inside omg_destroy for > 200 s)
  invalidate_programs(h);
  cudaStreamSynchronize(h->stream2);
  if (h->comm) omg_comm_destroy(h->comm);
  try { drain_totals(h, true); collect_profile(h); free_levels(h); } catch (...) {}
  for (auto& r : h->trace_events) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  h->trace_events.clear();
  h->d_scal.free(); h->d_partial.free(); h->d_epoch.free();
  for (int q = 0; q < 2; ++q) {
    if (h->h_stage[q]) cudaFreeHost(h->h_stage[q]);
    if (h->d_stage[q]) cudaFree(h->d_stage[q]);
    if (h->d_boxidx[q]) cudaFree(h->d_boxidx[q]);
    if (h->h_boxidx[q]) cudaFreeHost(h->h_boxidx[q]);
    if (h->stage_ev[q]) cudaEventDestroy(h->stage_ev[q]);
  }
  cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1);
  cudaEventDestroy(h->ev_fork); cudaEventDestroy(h->ev_join);
  cudaStreamDestroy(h->stream2);
  cudaStreamDestroy(h->stream);
  delete h;
  return 0;
}

const char* omg_last_error(const omg_handle* h) { return h ? h->err.c_str() : "invalid handle (no CUDA device?)"; }

int omg_set_tree(omg_handle* h, const omg_tree_desc* t) { OMG_TRY(h, set_tree(h, t)); }

int omg_set_params(omg_handle* h, const omg_params* p) {
  OMG_TRY(h, {
    if (p->lambda < 0) fail("set_lambda: lambda < 0 not allowed");
    if (h->have_params && std::memcmp(&h->params, p, sizeof(omg_params)) == 0) return 0;   // unchanged: keep cached programs
    std::memset(&h->params, 0, sizeof(omg_params));
    h->params.operator_type = p->operator_type; h->params.smoother_type = p->smoother_type; h->params.geometry_type = p->geometry_type;
    h->params.n_cycle_down = p->n_cycle_down; h->params.n_cycle_up = p->n_cycle_up; h->params.max_coarse_cycles = p->max_coarse_cycles;
    h->params.subtract_mean = p->subtract_mean; h->params.residual_coarse_abs = p->residual_coarse_abs;
    h->params.residual_coarse_rel = p->residual_coarse_rel; h->params.lambda = p->lambda;
    h->have_params = true; h->params_dirty = true;
  });
}

int omg_set_bc_const(omg_handle* h, int iv, int nb, int bc_type, double bc_value) {
  OMG_TRY(h, {
    if (iv < 1 || iv > OMG_MAXV || nb < 1 || nb > 6) fail("omg_set_bc_const: index out of range");
    if (bc_type != -10 && bc_type != -11 && bc_type != -12) fail("bc_to_gc: unknown boundary condition");
    if (h->bc_type[iv - 1][nb - 1] == bc_type && h->bc_value[iv - 1][nb - 1] == bc_value) return 0;
    h->bc_type[iv - 1][nb - 1] = bc_type; h->bc_value[iv - 1][nb - 1] = bc_value; h->params_dirty = true;
  });
}

int omg_set_bc_plane(omg_handle* h, int box_id, int nb, int bc_type, const double* plane) {
  OMG_TRY(h, {
    if (!h->have_tree) fail("no tree");
    if (box_id < 1 || box_id > h->n_boxes || h->loc[box_id] < 0) fail("omg_set_bc_plane: box not owned by this rank");
    if (nb < 1 || nb > 6) fail("omg_set_bc_plane: nb out of range");
    if (bc_type != -10 && bc_type != -11 && bc_type != -12) fail("bc_to_gc: unknown boundary condition");
    LevelHost& L = h->L(h->blvl[box_id - 1]);
    const size_t q = (size_t)h->loc[box_id] * 6 + (nb - 1);
    if (L.h_fkind[q] != FK_PHYS) fail("omg_set_bc_plane: not a physical boundary face");
    const size_t n2 = (size_t)L.nc * L.nc;
    if (L.h_farg[q] < 0) { L.h_farg[q] = L.n_bcpl++; L.h_farg_old[q] = L.h_farg[q]; L.h_bcpl.resize((size_t)L.n_bcpl * n2); }
    std::memcpy(&L.h_bcpl[(size_t)L.h_farg[q] * n2], plane, n2 * sizeof(double));
    L.h_ftype[q] = bc_type;
    L.meta_dirty = true;
  });
}

int omg_push_var(omg_handle* h, int iv, int n, const int32_t* ids, const double* const* ptrs) {
  OMG_TRY(h, push_pull_ids(h, iv, n, ids, ptrs, nullptr, 0, true));
}
int omg_pull_var(omg_handle* h, int iv, int n, const int32_t* ids, double* const* ptrs, int with_ghosts) {
  OMG_TRY(h, push_pull_ids(h, iv, n, ids, nullptr, ptrs, with_ghosts, false));
}
int omg_push_var_packed(omg_handle* h, int iv, int lvl_lo, int lvl_hi, const double* buf, int64_t n_doubles) {
  OMG_TRY(h, push_pull_packed(h, iv, lvl_lo, lvl_hi, buf, nullptr, n_doubles, 0, true));
}
int omg_pull_var_packed(omg_handle* h, int iv, int lvl_lo, int lvl_hi, double* buf, int64_t n_doubles, int with_ghosts) {
  OMG_TRY(h, push_pull_packed(h, iv, lvl_lo, lvl_hi, nullptr, buf, n_doubles, with_ghosts, false));
}
int64_t omg_packed_size(omg_handle* h, int lvl_lo, int lvl_hi) {
  if (!h || !h->have_tree) return -1;
  int64_t n = 0;
  for (int lvl = std::max(lvl_lo, h->lowest); lvl <= std::min(lvl_hi, h->highest); ++lvl) {
    LevelHost& L = h->L(lvl);
    n += (int64_t)(L.nc + 2) * (L.nc + 2) * (L.nc + 2) * L.nb;
  }
  return n;
}
// mg_copy_to_tree / mg_copy_from_tree[_gc] (coupling_amrvac/mod_multigrid_coupling.t:133-171,206-267): the application's
// grids <-> the tree, here host buffer <-> device tree in one pipelined pass (no host cc arrays in between)
int omg_copy_to_tree(omg_handle* h, int iv, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts, const double* buf, int64_t n_doubles) {
  OMG_TRY(h, push_pull_packed(h, iv, lvl_lo, lvl_hi, buf, nullptr, n_doubles, 0, true, with_ghosts ? 1 : 0, leaves_only != 0));
}
int omg_copy_from_tree(omg_handle* h, int iv, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts, double* buf, int64_t n_doubles) {
  OMG_TRY(h, push_pull_packed(h, iv, lvl_lo, lvl_hi, nullptr, buf, n_doubles, with_ghosts, false, with_ghosts ? 1 : 0, leaves_only != 0));
}
int64_t omg_copy_size(omg_handle* h, int lvl_lo, int lvl_hi, int leaves_only, int with_ghosts) {
  if (!h || !h->have_tree) return -1;
  int64_t n = 0;
  for (int lvl = std::max(lvl_lo, h->lowest); lvl <= std::min(lvl_hi, h->highest); ++lvl) {
    LevelHost& L = h->L(lvl);
    const int64_t edge = L.nc + (with_ghosts ? 2 : 0);
    n += edge * edge * edge * (int64_t)(leaves_only ? L.h_leaves.size() : (size_t)L.nb);
  }
  return n;
}
int omg_set_var_const(omg_handle* h, int iv, int lvl_lo, int lvl_hi, double value, int interior_only) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    if (!h->have_tree) fail("no tree: call omg_set_tree first");
    const int v = check_iv(h, iv);
    for (int lvl = std::max(lvl_lo, h->lowest); lvl <= std::min(lvl_hi, h->highest); ++lvl) {
      LevelHost& L = h->L(lvl);
      omgk::fill(L.d_var[v].p, L.d_var[v].n, value, h->stream);
      if (!interior_only && L.d_faces[v].n) omgk::fill(L.d_faces[v].p, L.d_faces[v].n, value, h->stream);
      if (!interior_only && v == IPHI && L.d_gpl.n) omgk::fill(L.d_gpl.p, L.d_gpl.n, value, h->stream);
    }
  });
}

int omg_fas_vcycle(omg_handle* h, int has_highest_lvl, int highest_lvl, int standalone, int want_res, double* max_res) {
  OMG_TRY(h, fas_vcycle(h, has_highest_lvl, highest_lvl, standalone, want_res, max_res));
}
int omg_fas_fmg(omg_handle* h, int have_guess, int want_res, double* max_res) {
  OMG_TRY(h, fas_fmg(h, have_guess, want_res, max_res));
}

int omg_fas_vcycles(omg_handle* h, int n, double* residuals) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    if (n < 0 || n > 256) fail("omg_fas_vcycles: n out of range (0..256)");
    check_p2p(h);
    sync_meta(h);
    check_supported(h);
    const int want = residuals ? 1 : 0;
    char key[64];
    std::snprintf(key, sizeof key, "v:%d:%d:%d:%d", 0, 0, 1, want);
    bool fresh;
    Program& P = get_program(h, key, fresh);
    if (fresh) { Builder B{h, P, {}}; B.vcycle(false, 0, true, want != 0, SCAL_MAXRES); B.flush(); }
    CK(cudaEventRecord(h->ev0, h->stream));
    for (int c = 0; c < n; ++c) {
      run_program(h, P);
      if (want) CK(cudaMemcpyAsync(h->d_scal.p + SCAL_RESHIST + c, h->d_scal.p + SCAL_MAXRES, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    }
    CK(cudaEventRecord(h->ev1, h->stream));
    if (want && n) CK(cudaMemcpyAsync(residuals, h->d_scal.p + SCAL_RESHIST, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    check_p2p(h);
    float ms = 0; CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1)); h->last_ms = ms;
    h->timer_acc[0] += ms * 1e-3;
  });
}

// Run one building block of the cycle on its own (parity tests compare each against the oracle):
// "smooth_boxes" (lvl, n_cycle)  src/m_multigrid.f90:404-424     "update_coarse" (lvl)  :347-384
// "correct_children" (lvl)  :387-402 (+ the ghost fill that follows it)   "coarse_solve"  :202-208
// "max_residual_lvl" (lvl) -> *out  :296-311
int omg_debug_step(omg_handle* h, const char* what, int lvl, int arg, double* out) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    check_p2p(h);
    sync_meta(h);
    check_supported(h);
    const std::string w(what);
    if (lvl < h->lowest || lvl > h->highest) fail("omg_debug_step: level out of range");
    Program P;
    for (int l = 0; l < OMG_NLVL; ++l) h->gz[l] = false;
    Builder B{h, P, {}};
    if (w == "smooth_boxes") B.smooth_boxes(lvl, arg);
    else if (w == "smooth_only") { for (int n = 1; n <= 2 * arg; ++n) B.emit(lvl, TOP_SMOOTH, S_SMOOTH, (n - 1) & 1); }   // timing probe: no ghost exchange
    else if (w == "halo_only") { for (int n = 1; n <= 2 * arg; ++n) B.gf(lvl); }                                          // timing probe
    else if (w == "update_coarse") { if (lvl <= h->lowest) fail("update_coarse: lvl <= lowest_lvl"); B.update_coarse(lvl); }
    else if (w == "correct_children") { if (lvl >= h->highest) fail("correct_children: lvl >= highest_lvl"); B.correct_children_onto(lvl + 1, 0); }   // (0: the ghost fill as a stand-alone exchange)
    else if (w == "coarse_solve") {
      const int sub = (h->params.smoother_type == 2) ? 2 : 1;
      B.pend.push_back({TOP_COARSE_SOLVE, B.li(h->lowest), (h->params.n_cycle_up + h->params.n_cycle_down) * sub, h->params.max_coarse_cycles});
    } else if (w == "max_residual_lvl") {
      B.step(S_CLEAR_SLOT, 0, SCAL_MAXRES); B.emit(lvl, TOP_RESID, S_RESID, SCAL_MAXRES);
      if (B.multi()) B.step(S_ALLREDUCE_MAX, 0, SCAL_MAXRES);
    }
    else fail("omg_debug_step: unknown step " + w);
    B.flush();
    const int ug = h->use_graph; h->use_graph = 0;
    try { run_program(h, P); } catch (...) { h->use_graph = ug; P.d_ops.free(); throw; }
    h->use_graph = ug;
    CK(cudaStreamSynchronize(h->stream));
    check_p2p(h);
    P.d_ops.free();
    if (out) CK(cudaMemcpy(out, h->d_scal.p + SCAL_MAXRES, sizeof(double), cudaMemcpyDeviceToHost));
  });
}

int omg_fill_ghost_cells_lvl(omg_handle* h, int lvl, int iv) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    check_p2p(h);
    sync_meta(h);
    const int v = check_iv(h, iv);
    if (lvl < h->lowest) fail("fill_ghost_cells_lvl: lvl < lowest_lvl");
    if (lvl > h->highest) fail("fill_ghost_cells_lvl: lvl > highest_lvl");
    LevelHost& L = h->L(lvl);
    if (v == IPHI) {
      if (L.xg.any) exchange_ghost(h, lvl, 0, h->stream);
      if (!L.h_rb.empty()) omgk::rb_fill(L.dev, h->L(lvl - 1).dev, h->pdev, 0, 0, h->stream);
    } else if (v >= 4) {
      // faces owned by other ranks (buffer_for_nb / buffer_for_fine_nb with this variable, src/m_ghost_cells.f90:142-168):
      // packed send/recv into a receive buffer, copied into the stored-face pool by fill_stored / used by rb_fill
      const double* rpl = nullptr;
      if (L.xg.any) {
        const LevelDev& Lc = lvl > h->lowest ? h->L(lvl - 1).dev : L.dev;
        if (L.xg.d_send.n < (size_t)L.xg.n_send * L.nc * L.nc) L.xg.d_send.alloc((size_t)L.xg.n_send * L.nc * L.nc);
        if (L.xg.d_recv.n < (size_t)L.xg.n_recv * L.nc * L.nc) L.xg.d_recv.alloc((size_t)L.xg.n_recv * L.nc * L.nc);
        if (L.xg.n_send) { omgk::pack_ghost(L.dev, Lc, h->pdev, L.xg.d_sdesc.p, L.xg.n_send, L.xg.d_send.p, 0, h->stream, v); h->launches++; }
        run_exchange(h, L.xg, L.xg.d_send.p, L.xg.d_recv.p, (size_t)L.nc * L.nc, h->stream);
        rpl = L.xg.d_recv.p;
      }
      omgk::fill_stored(L.dev, h->pdev, v, h->bc_type[v], h->bc_value[v], h->stream, rpl, L.xg.n_recv);
      if (!L.h_rb.empty()) omgk::rb_fill(L.dev, h->L(lvl - 1).dev, h->pdev, v, 0, h->stream, rpl);
    } else fail("fill_ghost_cells_lvl: ghost cells of rhs/old/res are not tracked on the device");
    h->launches++;
  });
}

int omg_restrict_lvl(omg_handle* h, int iv, int lvl) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    check_p2p(h);
    sync_meta(h);
    const int v = check_iv(h, iv);
    if (lvl <= h->lowest || lvl > h->highest) fail("cannot restrict lvl <= lowest_lvl");
    omgk::restrict2(h->L(lvl).dev, h->L(lvl - 1).dev, v, -1, h->stream);
    if (h->L(lvl).xr.any) exchange_restrict(h, lvl, v, -1, h->stream);
    h->launches++;
  });
}

int omg_prolong(omg_handle* h, int lvl, int iv, int iv_to, int add) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    sync_meta(h);
    const int v = check_iv(h, iv); const int vt = check_iv(h, iv_to);
    if (lvl == h->highest) fail("cannot prolong highest level");
    if (lvl < h->lowest) fail("cannot prolong below lowest level");
    if (!(v == IPHI || v >= 4)) fail("omg_prolong: source must be phi or an extra variable (ghost cells of rhs/old/res are not tracked)");
    check_p2p(h);
    if (h->L(lvl + 1).xp.any) {
      // prolong_set_buffer (src/m_prolong.f90:91-121) as a patch of coarse values, see do_pack_prolong_var
      LevelHost& Lf = h->L(lvl + 1);
      const int pw = Lf.nc / 2 + 2;
      if (Lf.xp.n_send) { omgk::pack_prolong_var(h->L(lvl).dev, h->pdev, Lf.xp.d_sdesc.p, Lf.xp.n_send, Lf.xp.d_send.p, Lf.nc, v, h->stream); h->launches++; }
      run_exchange(h, Lf.xp, Lf.xp.d_send.p, Lf.xp.d_recv.p, (size_t)pw * pw * pw, h->stream);
    }
    omgk::prolong_var(h->L(lvl + 1).dev, h->L(lvl).dev, h->pdev, v, vt, add, h->stream);
    h->launches++;
  });
}

int omg_apply_op(omg_handle* h, int i_out) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    sync_meta(h);
    check_supported(h);
    const int v = check_iv(h, i_out);
    for (LevelHost& L : h->levels) { omgk::apply_op(L.dev, h->pdev, v, h->stream); h->launches++; }
  });
}

// set_rhs of src/m_diffusion.f90:159-173: rhs <- f1 * phi + f2 * rhs on the owned leaves of levels 1..highest
int omg_diffusion_set_rhs(omg_handle* h, double f1, double f2) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    if (!h->have_tree) fail("no tree: call omg_set_tree first");
    sync_meta(h);
    for (int lvl = std::max(1, h->lowest); lvl <= h->highest; ++lvl) {
      LevelHost& L = h->L(lvl);
      if (L.h_leaves.empty()) continue;
      omgk::axpby_leaves(L.dev, IPHI, IRHS, f1, f2, h->stream); h->launches++;
    }
  });
}


// ---- free-space front end (include/omg_gpu.h; src/m_free_space.f90:36-214) ---------------------------------------------
int omg_free_space_kernel(omg_handle* h, int fft_lvl, const double r_min[3], int n_gauss, int n_range, const double* w,
                          const double* kx, const double* ky, const double* kz) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    if (!h->have_tree) fail("no tree: call omg_set_tree first");
    if (fft_lvl < h->lowest || fft_lvl > h->highest) fail("mg_poisson_free_3d: FFT level out of range");
    if (n_gauss < 1 || n_range < 1 || !w || !kx || !ky || !kz) fail("mg_poisson_free_3d: bad kernel");
    LevelHost& LF = h->L(fft_lvl);
    const int nc = LF.nc;
    int nbox[3] = {0, 0, 0};
    for (int id : h->lvl_ids[fft_lvl - h->lowest]) for (int d = 0; d < 3; ++d) nbox[d] = std::max(nbox[d], h->ix[(size_t)(id - 1) * 3 + d]);
    if ((size_t)nbox[0] * nbox[1] * nbox[2] != h->lvl_ids[fft_lvl - h->lowest].size()) fail("mg_poisson_free_3d: the FFT level is not fully refined");
    const int nx[3] = {nbox[0] * nc + 2, nbox[1] * nc + 2, nbox[2] * nc + 2};
    if (n_range < std::max(nx[0], std::max(nx[1], nx[2])) - 1) fail("mg_poisson_free_3d: kernel range smaller than the FFT grid");
    if (h->fs_plan) { omgfs::destroy(h->fs_plan); h->fs_plan = nullptr; }
    std::string err;
    h->fs_plan = omgfs::create(nx, err);
    if (!h->fs_plan) fail(err);
    if (!omgfs::set_kernel(h->fs_plan, n_gauss, n_range, w, kx, ky, kz, h->stream, err)) fail(err);
    h->fs_lvl = fft_lvl;
    for (int d = 0; d < 3; ++d) { h->fs_geom.r_min[d] = r_min[d] - 0.5 * LF.dev.dr[d]; h->fs_geom.inv_dr[d] = 1 / LF.dev.dr[d]; }
    std::vector<int> origin((size_t)LF.nb * 3);
    for (int b = 0; b < LF.nb; ++b) for (int d = 0; d < 3; ++d) origin[(size_t)b * 3 + d] = (h->ix[(size_t)(LF.ids[b] - 1) * 3 + d] - 1) * nc;
    h->fs_origin.upload(origin, h->stream);
    // every physical-boundary face of every level gets a Dirichlet plane (mg%bc(:, mg_iphi)%boundary_cond =>
    // ghost_cells_free_bc, :98-100, stored once by mg_phi_bc_store, :170)
    h->fs_faces.resize(h->levels.size()); h->fs_nfaces.assign(h->levels.size(), 0);
    for (size_t li = 0; li < h->levels.size(); ++li) {
      LevelHost& L = h->levels[li];
      const size_t n2 = (size_t)L.nc * L.nc;
      std::vector<omgfs::Face> faces;
      for (int b = 0; b < L.nb; ++b) for (int f = 0; f < 6; ++f) {
        const size_t q = (size_t)b * 6 + f;
        if (L.h_fkind[q] != FK_PHYS) continue;
        if (L.h_farg[q] < 0) { L.h_farg[q] = L.n_bcpl++; L.h_farg_old[q] = L.h_farg[q]; L.h_bcpl.resize((size_t)L.n_bcpl * n2, 0.0); }
        L.h_ftype[q] = -10;
        const int d = f >> 1, ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
        const int* ix = &h->ix[(size_t)(L.ids[b] - 1) * 3];
        faces.push_back({L.h_farg[q], f, r_min[ta] + (ix[ta] - 1) * (L.nc * L.dev.dr[ta]), r_min[tb] + (ix[tb] - 1) * (L.nc * L.dev.dr[tb])});
      }
      L.meta_dirty = true;
      h->fs_nfaces[li] = (int)faces.size();
      h->fs_faces[li].upload(faces, h->stream);
    }
    CK(cudaStreamSynchronize(h->stream));
  });
}

int omg_free_space_solve(omg_handle* h, double rhs_fac) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    check_p2p(h);
    if (!h->fs_plan) fail("mg_poisson_free_3d: no kernel (omg_free_space_kernel first)");
    sync_meta(h);
    LevelHost& LF = h->L(h->fs_lvl);
    cudaStream_t s = h->stream;
    int nx[3]; { int nbox[3] = {0, 0, 0}; for (int id : h->lvl_ids[h->fs_lvl - h->lowest]) for (int d = 0; d < 3; ++d) nbox[d] = std::max(nbox[d], h->ix[(size_t)(id - 1) * 3 + d]); for (int d = 0; d < 3; ++d) nx[d] = nbox[d] * LF.nc + 2; }
    double* grid = omgfs::grid(h->fs_plan);
    const size_t N = omgfs::grid_size(h->fs_plan);
    CK(cudaMemsetAsync(grid, 0, N * sizeof(double), s));
    omgfs::gather_rhs(LF.dev, h->fs_origin.p, rhs_fac, grid, nx, s); h->launches++;
    if (h->n_cpu > 1) { need_comm(h); try { omg_comm_allreduce_sum_vec(h->comm, grid, N, s); } catch (const std::exception& e) { fail(e.what()); } }
    std::string err;
    if (!omgfs::convolve(h->fs_plan, LF.dev.dr[0] * LF.dev.dr[1] * LF.dev.dr[2], s, err)) fail(err);
    h->launches += 5;
    for (size_t li = 0; li < h->levels.size(); ++li) {
      LevelHost& L = h->levels[li];
      if (!h->fs_nfaces[li]) continue;
      omgfs::bc_planes(L.dev, h->fs_faces[li].p, h->fs_nfaces[li], grid, nx, h->fs_geom, s); h->launches++;
      // the host copy follows (sync_meta re-uploads it whenever the level's metadata change; omg_get_bc_plane reads it)
      CK(cudaMemcpyAsync(L.h_bcpl.data(), L.d_bcpl.p, L.h_bcpl.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    omgfs::scatter_phi(LF.dev, h->fs_origin.p, grid, nx, s); h->launches++;
    // the reference copies the ghost layer along with the solution (cc(0:nc+1, ...), :176-179); here ghost values are
    // derived, except those of faces owned by other ranks, which need their exchange before the level is prolonged
    if (LF.xg.any) exchange_ghost(h, h->fs_lvl, 0, s);
    CK(cudaStreamSynchronize(s));
    check_p2p(h);
  });
}

int omg_get_bc_plane(omg_handle* h, int box_id, int nb, int* bc_type, double* plane) {
  OMG_TRY(h, {
    if (!h->have_tree) fail("no tree");
    if (box_id < 1 || box_id > h->n_boxes || h->loc[box_id] < 0) fail("omg_get_bc_plane: box not owned by this rank");
    if (nb < 1 || nb > 6) fail("omg_get_bc_plane: nb out of range");
    LevelHost& L = h->L(h->blvl[box_id - 1]);
    const size_t q = (size_t)h->loc[box_id] * 6 + (nb - 1);
    if (L.h_fkind[q] != FK_PHYS || L.h_farg[q] < 0) fail("omg_get_bc_plane: no stored plane on this face");
    const size_t n2 = (size_t)L.nc * L.nc;
    std::memcpy(plane, &L.h_bcpl[(size_t)L.h_farg[q] * n2], n2 * sizeof(double));
    if (bc_type) *bc_type = L.h_ftype[q];
  });
}

int omg_nccl_unique_id(char out[128]) { return omg_comm_unique_id(out); }
// After the communicator exists: map every distributed level's plane pool of every rank (collective, same level
// order on all ranks because xg.any comes from the replicated tree).  Call order contract: omg_set_tree, then
// omg_comm_init (mg_comm_init runs before mg_allocate_storage in the reference, but the device tree only exists
// after it; the host mirror / shim call omg_comm_init last).
void reset_epochs(omg_handle* h) {
  CK(cudaSetDevice(h->device));
  if (!h->d_epoch.p) h->d_epoch.alloc(OMG_NLVL);
  CK(cudaMemset(h->d_epoch.p, 0, OMG_NLVL * sizeof(unsigned long long)));
  h->params_dirty = true;                    // LevelDev copies carry gpl_sel
}
// map every distributed level's plane pool of every rank (NCCL back end; collective, same level order on all ranks)
void share_pools(omg_handle* h) {
  const int nranks = h->n_cpu;
  for (LevelHost& L : h->levels) L.xg.shared = false;
  if (!h->comm || !h->have_tree || !omg_comm_p2p(h->comm) || !omg_comm_graph_safe(h->comm)) return;
  std::vector<void*> ptrs(nranks, nullptr);
  for (LevelHost& L : h->levels) {
    if (!L.xg.any) continue;
    try { omg_comm_share(h->comm, L.d_gpl.p, ptrs.data()); } catch (const std::exception& e) { fail(e.what()); }
    if (!omg_comm_p2p(h->comm)) break;                       // IPC unavailable: every rank falls back to send/recv
    std::vector<double*> pb(nranks);
    for (int r = 0; r < nranks; ++r) pb[r] = static_cast<double*>(ptrs[r]);
    L.xg.d_peer_base.upload(pb, h->stream);
    CK(cudaStreamSynchronize(h->stream));
    L.xg.shared = true;
  }
  if (!omg_comm_p2p(h->comm)) for (LevelHost& L : h->levels) L.xg.shared = false;
}
int omg_comm_init(omg_handle* h, const char uid[128], int rank, int nranks) {
  OMG_TRY(h, {
    invalidate_programs(h);                  // graphs that captured the old communicator's operations go first
    if (h->comm) { omg_comm_destroy(h->comm); h->comm = nullptr; }
    try { h->comm = omg_comm_create_nccl(uid, rank, nranks, h->device); } catch (const std::exception& e) { fail(e.what()); }
    reset_epochs(h);
    share_pools(h);
    invalidate_programs(h);
  });
}

// ---- on-device regrid / re-partition (include/omg_gpu.h: omg_update_tree) ----------------------------------------------
namespace {
struct OldLevel { int lvl = 0, nc = 0; std::vector<int> ids; DevBuf<double> var[OMG_MAXV], faces[OMG_MAXV]; };

void update_tree(omg_handle* h, const omg_tree_desc* t) {
  if (!h->have_tree) { set_tree(h, t); return; }
  CK(cudaSetDevice(h->device));
  if (t->box_size != h->box_size || t->n_extra_vars != h->n_extra || t->n_cpu != h->n_cpu || t->my_rank != h->my_rank)
    fail("omg_update_tree: box size, number of variables and ranks must stay the same");
  if (h->n_cpu > 1) need_comm(h);
  CK(cudaStreamSynchronize(h->stream));
  check_p2p(h);
  const int me = h->my_rank, nvar = h->nvar;
  // the peers' mappings of the old plane pools must be closed before the pools are freed: everybody closes, then a barrier
  if (h->comm && omg_comm_graph_safe(h->comm)) {
    invalidate_programs(h);
    try { omg_comm_unshare(h->comm); } catch (const std::exception& e) { fail(e.what()); }
  }
  // 1. keep the old variables out of free_levels' reach
  std::vector<OldLevel> old(h->levels.size());
  for (size_t li = 0; li < h->levels.size(); ++li) {
    LevelHost& L = h->levels[li];
    old[li].lvl = L.lvl; old[li].nc = L.nc; old[li].ids = L.ids;
    for (int v = 0; v < OMG_MAXV; ++v) { std::swap(old[li].var[v], L.d_var[v]); std::swap(old[li].faces[v], L.d_faces[v]); }
  }
  const std::vector<int> old_rank = h->rank, old_lvl = h->blvl;
  const int old_n = h->n_boxes, old_lowest = h->lowest;
  auto free_old = [&]() { for (OldLevel& O : old) for (int v = 0; v < OMG_MAXV; ++v) { O.var[v].free(); O.faces[v].free(); } };
  try {
    // 2. the new tree: fresh, zero-filled storage (mg_allocate_storage, src/m_allocate_storage.f90:72-76)
    set_tree(h, t);
    cudaStream_t s = h->stream;
    // 3. boxes that exist in both trees on the same level keep their data
    for (LevelHost& L : h->levels) {
      const int oi = L.lvl - old_lowest;
      if (oi < 0 || oi >= (int)old.size() || old[oi].nc != L.nc) continue;
      OldLevel& O = old[oi];
      std::map<int, int> old_loc;
      for (size_t b = 0; b < O.ids.size(); ++b) old_loc[O.ids[b]] = (int)b;
      const size_t bs = (size_t)L.dev.bs, fs = (size_t)6 * L.nc * L.nc;
      // (a) stayed on this rank: device-to-device
      std::vector<int> pairs;
      for (int b = 0; b < L.nb; ++b) { auto it = old_loc.find(L.ids[b]); if (it != old_loc.end()) { pairs.push_back(it->second); pairs.push_back(b); } }
      if (!pairs.empty()) {
        DevBuf<int> d_pairs; d_pairs.upload(pairs, s);
        for (int v = 0; v < nvar; ++v) {
          omgk::copy_boxes(L.d_var[v].p, O.var[v].p, d_pairs.p, (int)pairs.size() / 2, bs, s); h->launches++;
          if (v >= 4 && O.faces[v].p && L.d_faces[v].p) { omgk::copy_boxes(L.d_faces[v].p, O.faces[v].p, d_pairs.p, (int)pairs.size() / 2, fs, s); h->launches++; }
        }
        CK(cudaStreamSynchronize(s));
        d_pairs.free();
      }
      // (b) changed owner: old owner -> new owner, ids ascending on both sides (the tree is replicated)
      if (h->n_cpu > 1) {
        std::map<int, std::vector<int>> send_to, recv_from;       // peer -> box ids
        for (int id : h->lvl_ids[L.lvl - h->lowest]) {
          if (id > old_n || old_lvl[id - 1] != L.lvl) continue;
          const int ro = old_rank[id - 1], rn = h->rank[id - 1];
          if (ro == rn) continue;
          if (ro == me) send_to[rn].push_back(id);
          if (rn == me) recv_from[ro].push_back(id);
        }
        // every rank takes part in the exchange of a level on which ANY box moved (same decision everywhere)
        bool any = false;
        for (int id : h->lvl_ids[L.lvl - h->lowest]) if (id <= old_n && old_lvl[id - 1] == L.lvl && old_rank[id - 1] != h->rank[id - 1]) any = true;
        if (any) {
          const size_t item = (size_t)nvar * bs + (size_t)h->n_extra * fs;
          std::vector<int> peers;
          for (auto& kv : send_to) peers.push_back(kv.first);
          for (auto& kv : recv_from) if (!send_to.count(kv.first)) peers.push_back(kv.first);
          std::sort(peers.begin(), peers.end());
          size_t ns = 0, nr = 0;
          for (auto& kv : send_to) ns += kv.second.size();
          for (auto& kv : recv_from) nr += kv.second.size();
          DevBuf<double> d_send, d_recv; d_send.alloc(ns * item); d_recv.alloc(nr * item);
          std::vector<OmgXfer> segs; std::vector<int> spairs, rpairs;     // (box index in its array, slot in the buffer)
          size_t so = 0, ro = 0;
          for (int p : peers) {
            OmgXfer x{p, d_send.p + so * item, 0, d_recv.p + ro * item, 0};
            if (send_to.count(p)) { for (int id : send_to[p]) { spairs.push_back(old_loc.at(id)); spairs.push_back((int)so++); } x.send_n = send_to[p].size() * item; }
            if (recv_from.count(p)) { for (int id : recv_from[p]) { rpairs.push_back((int)ro++); rpairs.push_back(h->loc[id]); } x.recv_n = recv_from[p].size() * item; }
            segs.push_back(x);
          }
          DevBuf<int> d_sp, d_rp; d_sp.upload(spairs, s); d_rp.upload(rpairs, s);
          size_t off = 0;
          for (int v = 0; v < nvar; ++v) {       // buffer layout per box: var 0 .. nvar-1, then the stored faces of the extra variables
            if (ns) { omgk::copy_boxes_strided(d_send.p + off, item, O.var[v].p, bs, d_sp.p, (int)ns, bs, s); h->launches++; }
            off += bs;
          }
          for (int v = 4; v < nvar; ++v) { if (ns && O.faces[v].p) { omgk::copy_boxes_strided(d_send.p + off, item, O.faces[v].p, fs, d_sp.p, (int)ns, fs, s); h->launches++; } off += fs; }
          try { omg_comm_exchange(h->comm, segs.data(), (int)segs.size(), s); } catch (const std::exception& e) { fail(e.what()); }
          off = 0;
          for (int v = 0; v < nvar; ++v) {
            if (nr) { omgk::copy_boxes_strided(L.d_var[v].p, bs, d_recv.p + off, item, d_rp.p, (int)nr, bs, s); h->launches++; }
            off += bs;
          }
          for (int v = 4; v < nvar; ++v) { if (nr && L.d_faces[v].p) { omgk::copy_boxes_strided(L.d_faces[v].p, fs, d_recv.p + off, item, d_rp.p, (int)nr, fs, s); h->launches++; } off += fs; }
          CK(cudaStreamSynchronize(s));
          d_send.free(); d_recv.free(); d_sp.free(); d_rp.free();
        }
      }
    }
  } catch (...) { free_old(); throw; }
  free_old();
  // 4. the new plane pools are mapped again (NCCL back end; emulated ranks call omg_comm_init_local once more)
  if (h->comm) { reset_epochs(h); share_pools(h); }
  invalidate_programs(h);
}
}  // namespace

int omg_update_tree(omg_handle* h, const omg_tree_desc* t) { OMG_TRY(h, update_tree(h, t)); }

int omg_comm_init_local(omg_handle* const* handles, int nranks) {
  if (!handles || nranks < 1) return 1;
  omg_handle* h0 = handles[0];
  OMG_TRY(h0, {
    omg_local_group* g = omg_local_group_create(nranks);
    for (int r = 0; r < nranks; ++r) {
      omg_handle* h = handles[r];
      if (h->comm) { omg_comm_destroy(h->comm); h->comm = nullptr; }
      try { h->comm = omg_comm_create_local(g, r, h->device); } catch (const std::exception& e) { fail(e.what()); }
      invalidate_programs(h);
      reset_epochs(h);
    }
    // emulated ranks live in one process: the "peer mappings" are the other handles' pointers themselves
    for (int r = 0; r < nranks; ++r) {
      omg_handle* h = handles[r];
      if (!h->have_tree) continue;
      CK(cudaSetDevice(h->device));
      for (size_t l = 0; l < h->levels.size(); ++l) {
        LevelHost& L = h->levels[l];
        L.xg.shared = false;
        if (!L.xg.any) continue;
        std::vector<double*> pb(nranks, nullptr);
        bool ok = true;
        for (int q = 0; q < nranks; ++q) {
          omg_handle* hq = handles[q];
          if (!hq->have_tree || hq->levels.size() != h->levels.size() || hq->device != h->device) { ok = false; break; }
          pb[q] = hq->levels[l].d_gpl.p;
        }
        if (!ok) continue;
        L.xg.d_peer_base.upload(pb, h->stream);
        CK(cudaStreamSynchronize(h->stream));
        L.xg.shared = true;
      }
    }
  });
}
/* CPU-only view of the static exchange plan (no device needed): keys of the items rank `me` sends to / receives
 * from `peer` in phase (0 ghost, 1 restrict, 2 prolong) on level lvl, in message order. */
int omg_exchange_plan(const omg_tree_desc* t, int me, int phase, int lvl, int peer, int64_t* send_keys, int64_t* recv_keys,
                      int cap, int* n_send, int* n_recv) {
  if (!t || phase < 0 || phase > 2) return 1;
  omgp::LevelPlan pl = omgp::build_plan(*t, me, phase, lvl);
  int ns = 0, nr = 0;
  for (const omgp::PeerPlan& pp : pl.peers) {
    if (pp.peer != peer) continue;
    for (const omgp::Item& it : pp.send) { if (send_keys && ns < cap) send_keys[ns] = it.key * 2 + it.kind; ++ns; }
    for (const omgp::Item& it : pp.recv) { if (recv_keys && nr < cap) recv_keys[nr] = it.key * 2 + it.kind; ++nr; }
  }
  if (n_send) *n_send = ns;
  if (n_recv) *n_recv = nr;
  return 0;
}

int omg_sync(omg_handle* h) { OMG_TRY(h, { CK(cudaSetDevice(h->device)); CK(cudaStreamSynchronize(h->stream)); check_p2p(h); }); }

// mg_timers_show (src/m_data_structures.f90:562-581): the seven timers mg_add_timers registers
// (src/m_multigrid.f90:72-81), accumulated seconds on this rank.  The two totals are always measured (CUDA events
// around each mg_fas_vcycle / mg_fas_fmg); the five phase timers only accumulate while option "profile" is set
// (a cycle replayed as one CUDA graph has no phase boundaries to time).  Call with names == seconds == NULL to get n.
int omg_timers(omg_handle* h, int* n, const char** names, double* seconds) {
  static const char* kNames[7] = {"mg total V-cycle", "mg total FMG cycle", "mg smoother", "mg smoother g.c.",
                                  "mg coarse", "mg correct", "mg update coarse"};
  OMG_TRY(h, {
    if (n) *n = 7;
    if (!names && !seconds) return 0;
    CK(cudaSetDevice(h->device));
    drain_totals(h, true);
    collect_profile(h);
    for (int t = 0; t < 7; ++t) { if (names) names[t] = kNames[t]; if (seconds) seconds[t] = h->timer_acc[t]; }
  });
}

int64_t omg_stat(omg_handle* h, const char* key) {
  if (!h) return -1;
  const std::string k(key);
  if (k == "launches") return h->launches;
  if (k == "fast_launches") return h->fast_launches;         // level kernels on the TMA-staged path (omg_fast.cu)
  if (k == "general_launches") return h->general_launches;   // level kernels on the general path (omg_kernels.cu)
  if (k == "fast_cells") return h->fast_cells;               // the same two, weighted by the cells of the level of each launch
  if (k == "general_cells") return h->general_cells;
  if (k == "graph_launches") return h->graph_launches;
  if (k == "bytes_h2d") return h->bytes_h2d;
  if (k == "bytes_d2h") return h->bytes_d2h;
  if (k == "exchanges") return h->exchanges;
  if (k == "fused_smooths") return h->fused_smooths;     // half-sweeps that exchanged their partition faces in-kernel
  if (k == "fused_prolongs") return h->fused_prolongs;   // prolongations that did
  if (k == "puts") return h->puts;                       // ghost phases sent as peer-mapped puts
  if (k == "coarse_iters") {
    double v = 0; cudaSetDevice(h->device); cudaStreamSynchronize(h->stream);
    if (h->d_scal.p) cudaMemcpy(&v, h->d_scal.p + SCAL_COARSE_ITERS, sizeof(double), cudaMemcpyDeviceToHost);
    return (int64_t)v;
  }
  if (k == "n_tail_levels") { int64_t c = 0; for (LevelHost& L : h->levels) c += L.tail; return c; }
  if (k == "device_bytes") { int64_t c = 0; for (LevelHost& L : h->levels) for (int v = 0; v < OMG_MAXV; ++v) c += (int64_t)(L.d_var[v].n + L.d_faces[v].n) * 8; return c; }
  return -1;
}

int omg_set_option(omg_handle* h, const char* key, int value) {
  OMG_TRY(h, {
    const std::string k(key);
    if (k == "use_graph") h->use_graph = value;
    else if (k == "fast_kernels") h->fast_kernels = value;
    else if (k == "graph_multi") h->graph_multi = value;
    else if (k == "tail_smem") h->tail_smem = value;
    else if (k == "tail_simple") { h->tail_simple = value; h->params_dirty = true; }
    else if (k == "overlap") h->overlap = value;
    else if (k == "pdl") h->pdl = value;
    else if (k == "ring") h->ring = value;
    else if (k == "coeff_math") { h->coeff_math = value; h->params_dirty = true; }
    else if (k == "warp_coarse") { h->warp_coarse = value; h->params_dirty = true; }
    else if (k == "ring_fused") h->ring_fused = value;
    else if (k == "ring_order") h->ring_order = value;   // fused levels on the ring kernel: 0 partition-face boxes first, 1 last, 2 natural order
    else if (k == "store_fine_residual") h->store_fine_residual = value;
    else if (k == "trace") { h->trace = value; for (auto& r : h->trace_events) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); } h->trace_events.clear(); }
    else if (k == "p2p") { h->p2p = value; h->params_dirty = true; }
    else if (k == "fused") { h->fused = value; h->params_dirty = true; }
    else if (k == "fused_prolong") { h->fused_prolong = value; h->params_dirty = true; }
    else if (k == "put_fence") h->put_fence = value;
    else if (k == "bnd_first_boxes") h->bnd_first_boxes = value;
    else if (k == "profile") { collect_profile(h); h->profile = value; if (value) h->prof_acc.clear(); }
    else if (k == "tail_cells") { h->tail_cells = value; if (h->have_tree) compute_tail(h); }
    else if (k == "tail_cluster") { h->tail_cluster = value; if (h->have_tree) compute_tail(h); h->params_dirty = true; }
    else if (k == "tail_cluster_cells") { h->tail_cluster_cells = value; if (h->have_tree) compute_tail(h); h->params_dirty = true; }
    else if (k == "tail_cluster_always") h->tail_cluster_always = value;
    else fail("unknown option " + k);
    invalidate_programs(h);
  });
}

// "cycles": elapsed ms of the last omg_fas_vcycles call.  With option "profile" set, per finest-level
// kernel class: "smooth", "resid_restrict", "corr_prolong", "resid" -> average ms per launch, and
// "n_<class>" -> number of launches measured.
double omg_last_ms(omg_handle* h, const char* what) {
  if (!h) return -1.0;
  const std::string k(what);
  if (k == "cycles") return h->last_ms;
  try { collect_profile(h); } catch (...) { return -1.0; }
  std::string name = k; bool count = false;
  if (k.rfind("n_", 0) == 0) { name = k.substr(2); count = true; }
  int kind = -1;
  if (name == "smooth") kind = S_SMOOTH; else if (name == "resid_restrict") kind = S_RESID_RESTRICT;
  else if (name == "corr_prolong") kind = S_CORR_PROLONG; else if (name == "resid") kind = S_RESID;
  else if (name == "restrict") kind = S_RESTRICT;
  auto it = h->prof_acc.find(kind);
  if (it == h->prof_acc.end() || it->second.second == 0) return count ? 0.0 : -1.0;
  return count ? (double)it->second.second : it->second.first / (double)it->second.second;
}
// Steps recorded since option "trace" was set: rec[i] = {step kind, level, a, b}, ms[i] = elapsed time of the step on the
// cycle's stream (steps that fork to the exchange stream show their issue cost; the join shows the wait).  At most cap
// records are written, *n receives the number available; the record list is cleared.
int omg_trace(omg_handle* h, int cap, int* rec, double* ms, int* n) {
  OMG_TRY(h, {
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    int i = 0;
    for (auto& r : h->trace_events) {
      if (i < cap && rec && ms) {
        float t = 0; CK(cudaEventElapsedTime(&t, r.e0, r.e1));
        rec[4 * i] = r.kind; rec[4 * i + 1] = r.l; rec[4 * i + 2] = r.a; rec[4 * i + 3] = r.b; ms[i] = t;
      }
      cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
      ++i;
    }
    if (n) *n = i;
    h->trace_events.clear();
  });
}
const char* omg_step_name(int kind) {
  static const char* names[] = {"smooth", "resid", "restrict", "crhs", "oldpl", "corr_prolong", "rb_fill", "copy_old", "coarse_solve",
                                "tail", "zero_phi", "submean", "clear_slot", "copy_slot", "halo", "xrestrict", "xprolong", "allreduce_max",
                                "resid_restrict", "halo_fork", "halo_join", "put_wait"};
  return (kind >= 0 && kind < (int)(sizeof(names) / sizeof(names[0]))) ? names[kind] : "?";
}
void* omg_stream(omg_handle* h) { return h ? (void*)h->stream : nullptr; }

}  // extern "C"
