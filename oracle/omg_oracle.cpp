// oracle/omg_oracle.cpp
//
// TEST INFRASTRUCTURE ONLY.  CPU restatement (NDIM = 3) of the octree-mg
// multigrid hot path, used as the checker for the CUDA implementation.  Only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load this library; the product (octree-mg_b200/) never
// links to it or calls it.
//
// PARITY STATUS: "parity unpinned" by golden vectors -- the reference ships no
// golden vectors / expected outputs and cannot be compiled in
// this image (Fortran + MPI, no compiler).  The restatement is
// pinned instead by (i) the reference tests' own pass criterion (manufactured
// solution, error -> rounding), (ii) rank-count invariance, (iii) GSRB
// box-size invariance, (iv) an independent global-array numpy restatement
// (oracle/global_array_mg.py) that agrees bitwise.  See tests/test_oracle_*.py.
//
// Every routine cites the reference file:line it restates.  Storage follows the
// reference: per box cc(0:nc+1, 0:nc+1, 0:nc+1, nvar), i fastest, variable
// slowest; box ids are 1-based; neighbors > 0 id, 0 = refinement boundary,
// < 0 physical boundary.  Build: g++ -O2 -ffp-contract=off (no FMA, no
// reassociation) to mimic gfortran -O2 on x86-64 (makerules.make:9-10).
//
// Multi-rank runs are emulated in-process: every "do i over my_ids" loop of the
// reference becomes a loop over emulated ranks, and sort_and_transfer_buffers
// (src/m_communication.f90:37-88) is restated with per-(src,dst) buffers that
// are key-sorted and then consumed sequentially by the receiver, exactly like
// the MPI code.  With OpenMP the ranks run as threads (CPU baseline).

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---- constants: src/m_data_structures.f90:8-87 ---------------------------
constexpr int LVL_LO = -20, LVL_HI = 20, NLVL = LVL_HI - LVL_LO + 1;
constexpr int OP_LAPLACIAN = 1, OP_VLAPLACIAN = 2, OP_HELMHOLTZ = 3,
              OP_VHELMHOLTZ = 4, OP_AHELMHOLTZ = 5;
constexpr int SM_GS = 1, SM_GSRB = 2;
constexpr int NUM_VARS = 4, MAX_NUM_VARS = 10;
// 0-based variable slots (Fortran mg_iphi=1 ... mg_iveps=5)
constexpr int IPHI = 0, IRHS = 1, IOLD = 2, IRES = 3, IVEPS = 4;
constexpr int BC_DIRICHLET = -10, BC_NEUMANN = -11, BC_CONTINUOUS = -12;
constexpr int NO_BOX = 0, PHYSICAL_BOUNDARY = -1;

// ---- topology tables: src/m_data_structures.f90:150-191 -------------------
// children are numbered 1..8, neighbours 1..6 as in the reference; the tables
// below are indexed 0-based and hold 1-based values where the reference does.
const int child_dix[8][3] = {{0,0,0},{1,0,0},{0,1,0},{1,1,0},
                             {0,0,1},{1,0,1},{0,1,1},{1,1,1}};
const int child_rev[3][8] = {{2,1,4,3,6,5,8,7},{3,4,1,2,7,8,5,6},{5,6,7,8,1,2,3,4}};
const int child_adj_nb[6][4] = {{1,3,5,7},{2,4,6,8},{1,2,5,6},{3,4,7,8},{1,2,3,4},{5,6,7,8}};
const bool neighb_low[6] = {true,false,true,false,true,false};
const int neighb_high_pm[6] = {-1,1,-1,1,-1,1};
const int neighb_rev[6] = {2,1,4,3,6,5};   // 1-based
const int neighb_dim[6] = {0,0,1,1,2,2};   // 0-based dimension

inline bool child_low(int d, int c1) { return child_dix[c1-1][d] == 0; }

typedef void (*bc_callback_t)(const double* r_min, const double* dr, int lvl, int nc,
                              int iv, int nb, int* bc_type, double* bc, void* user);

struct Box {
  int rank = 0, id = 0, lvl = 0, ix[3] = {0,0,0}, parent = 0;
  int children[8] = {0,0,0,0,0,0,0,0};
  int neighbors[6] = {0,0,0,0,0,0};
  double r_min[3] = {0,0,0}, dr[3] = {0,0,0};
  int nc = 0;                    // box size of its level
  std::vector<double> cc;        // (nc+2)^3 * nvar
};

struct Lvl {
  std::vector<int> leaves, parents, ref_bnds, ids;
  // per emulated rank
  std::vector<std::vector<int>> my_leaves, my_parents, my_ref_bnds, my_ids;
};

struct Bc {
  int bc_type = BC_DIRICHLET;
  double bc_value = 0.0;
  bc_callback_t boundary_cond = nullptr;
  void* user = nullptr;
};

struct Buf {           // one per ordered (src rank, dst rank) pair
  std::vector<double> send;
  std::vector<int> ix;
  size_t i_recv = 0;   // consumer cursor (owned by the dst rank)
};

struct Timer { std::string name; double t = 0, t0 = 0; };
struct FreeBc;                     // state of the free-space front end (free_space.inc)

struct Mg {
  bool tree_created = false, is_allocated = false;
  int n_extra_vars = 0;
  int n_cpu = 1;
  int box_size = -1, highest_lvl = -1, lowest_lvl = -1, first_normal_lvl = -1;
  int n_boxes = 0;
  int box_size_lvl[NLVL];
  int domain_size_lvl[NLVL][3];
  double dr[NLVL][3];
  double r_min[3] = {0,0,0};
  Lvl lvls[NLVL];
  std::vector<Box> boxes;          // boxes[0] unused (ids are 1-based)
  std::vector<Buf> buf;            // n_cpu * n_cpu, index src*n_cpu + dst
  bool phi_bc_data_stored = false;
  bool periodic[3] = {false,false,false};
  Bc bc[6][MAX_NUM_VARS];
  int operator_type = OP_LAPLACIAN;
  bool subtract_mean = false;
  int smoother_type = SM_GS;
  int n_smoother_substeps = 1;
  int n_cycle_down = 2, n_cycle_up = 2, max_coarse_cycles = 1000;
  int coarsest_grid[3] = {2,2,2};
  double residual_coarse_abs = 1e-8, residual_coarse_rel = 1e-8;
  double helmholtz_lambda = 0.0;   // module globals of m_helmholtz / m_vhelmholtz / m_ahelmholtz
  bool methods_set = false;
  int last_coarse_iters = 0;
  Timer timers[8];
  std::string err;
  FreeBc* free_bc = nullptr;

  int nvar() const { return NUM_VARS + n_extra_vars; }
  Lvl& L(int lvl) { return lvls[lvl - LVL_LO]; }
  int& bsl(int lvl) { return box_size_lvl[lvl - LVL_LO]; }
  double* drl(int lvl) { return dr[lvl - LVL_LO]; }
  Buf& B(int src, int dst) { return buf[(size_t)src * n_cpu + dst]; }
};

enum { T_VCYCLE, T_FMG, T_SMOOTH, T_SMOOTH_GC, T_COARSE, T_CORRECT, T_UPDATE_COARSE, N_TIMERS };

double wtime() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}
void tstart(Mg& mg, int t) { mg.timers[t].t0 = wtime(); }
void tend(Mg& mg, int t) { mg.timers[t].t += wtime() - mg.timers[t].t0; }

struct OracleError { std::string msg; };
[[noreturn]] void fail(const char* msg) { throw OracleError{msg}; }

// index into cc(0:nc+1, 0:nc+1, 0:nc+1, v)
inline size_t cix(int n2, int i, int j, int k, int v) {
  return (((size_t)v * n2 + k) * n2 + j) * n2 + i;
}

// ---- src/m_data_structures.f90:430-467 ------------------------------------
inline bool has_children(const Box& b) { return b.children[0] != NO_BOX; }
inline int ix_to_ichild(const int ix[3]) {
  return 8 - 4 * (ix[2] & 1) - 2 * (ix[1] & 1) - (ix[0] & 1);
}
inline void get_child_offset(const Mg& mg, int id, int off[3]) {
  const Box& b = mg.boxes[id];
  if (b.lvl <= mg.first_normal_lvl) { off[0] = off[1] = off[2] = 0; }
  else for (int d = 0; d < 3; ++d) off[d] = ((b.ix[d] - 1) & 1) * (mg.box_size >> 1);
}

// ===========================================================================
// Tree construction: src/m_build_tree.f90
// ===========================================================================

// src/m_build_tree.f90:258-293
void set_leaves_parents(Mg& mg, int lvl) {
  Lvl& L = mg.L(lvl);
  L.parents.clear(); L.leaves.clear();
  for (int id : L.ids) (has_children(mg.boxes[id]) ? L.parents : L.leaves).push_back(id);
}

// src/m_build_tree.f90:187-216
void set_next_level_ids(Mg& mg, int lvl) {
  Lvl& L = mg.L(lvl); Lvl& N = mg.L(lvl + 1);
  N.ids.clear();
  if (mg.bsl(lvl + 1) == mg.bsl(lvl)) {
    for (int id : L.parents) for (int c = 0; c < 8; ++c) N.ids.push_back(mg.boxes[id].children[c]);
  } else {
    for (int id : L.parents) N.ids.push_back(mg.boxes[id].children[0]);
  }
}

// src/m_build_tree.f90:236-255
int find_neighb(const Mg& mg, int id, int nb1) {
  int p_id = mg.boxes[id].parent;
  int c_ix = ix_to_ichild(mg.boxes[id].ix);
  int d = neighb_dim[nb1 - 1];
  if (child_low(d, c_ix) == neighb_low[nb1 - 1]) p_id = mg.boxes[p_id].neighbors[nb1 - 1];
  if (p_id <= NO_BOX) return p_id;   // (reference indexes boxes(p_id) only when > 0 in valid trees)
  return mg.boxes[p_id].children[child_rev[d][c_ix - 1] - 1];
}

// src/m_build_tree.f90:218-233 and :176-185
void set_neighbors_lvl(Mg& mg, int lvl) {
  for (int id : mg.L(lvl).ids) {
    for (int nb = 1; nb <= 6; ++nb) {
      if (mg.boxes[id].neighbors[nb - 1] == NO_BOX) {
        int nb_id = find_neighb(mg, id, nb);
        if (nb_id > NO_BOX) {
          mg.boxes[id].neighbors[nb - 1] = nb_id;
          mg.boxes[nb_id].neighbors[neighb_rev[nb - 1] - 1] = id;
        }
      }
    }
  }
}

// src/m_build_tree.f90:296-328
void set_refinement_boundaries(Mg& mg, int lvl) {
  Lvl& L = mg.L(lvl);
  L.ref_bnds.clear();
  if (L.parents.empty()) return;
  for (int id : L.leaves) {
    for (int nb = 0; nb < 6; ++nb) {
      int nb_id = mg.boxes[id].neighbors[nb];
      if (nb_id > NO_BOX && has_children(mg.boxes[nb_id])) { L.ref_bnds.push_back(id); break; }
    }
  }
}

// src/m_build_tree.f90:330-367
void add_children(Mg& mg, int id) {
  if (mg.n_boxes + 8 > (int)mg.boxes.size() - 1) mg.boxes.resize(mg.boxes.size() + 8192);
  int c_ids[8];
  for (int i = 0; i < 8; ++i) c_ids[i] = mg.n_boxes + 1 + i;
  mg.n_boxes += 8;
  int lvl = mg.boxes[id].lvl + 1;
  for (int i = 0; i < 8; ++i) mg.boxes[id].children[i] = c_ids[i];
  for (int i = 0; i < 8; ++i) {
    Box& c = mg.boxes[c_ids[i]];
    const Box& p = mg.boxes[id];
    c.rank = p.rank; c.id = c_ids[i]; c.lvl = lvl; c.parent = id;
    for (int d = 0; d < 3; ++d) c.ix[d] = 2 * p.ix[d] - 1 + child_dix[i][d];
    for (int n = 0; n < 8; ++n) c.children[n] = NO_BOX;
    for (int n = 0; n < 6; ++n) c.neighbors[n] = NO_BOX;
    for (int d = 0; d < 3; ++d) {
      c.r_min[d] = p.r_min[d] + mg.drl(lvl)[d] * child_dix[i][d] * mg.box_size;
      c.dr[d] = mg.drl(lvl)[d];
    }
    c.nc = mg.bsl(lvl);
  }
  for (int nb = 0; nb < 6; ++nb) {
    int pn = mg.boxes[id].neighbors[nb];
    if (pn < NO_BOX)
      for (int n = 0; n < 4; ++n) mg.boxes[c_ids[child_adj_nb[nb][n] - 1]].neighbors[nb] = pn;
  }
}

// src/m_build_tree.f90:369-393
void add_single_child(Mg& mg, int id, int n_boxes_lvl) {
  int c_id = mg.n_boxes + 1;
  mg.n_boxes += 1;
  Box& p = mg.boxes[id];
  Box& c = mg.boxes[c_id];
  p.children[0] = c_id;
  int lvl = p.lvl + 1;
  c.rank = p.rank; c.id = c_id; c.lvl = lvl; c.parent = id;
  for (int d = 0; d < 3; ++d) c.ix[d] = p.ix[d];
  for (int n = 0; n < 8; ++n) c.children[n] = NO_BOX;
  for (int n = 0; n < 6; ++n)
    c.neighbors[n] = (p.neighbors[n] == PHYSICAL_BOUNDARY) ? p.neighbors[n] : p.neighbors[n] + n_boxes_lvl;
  for (int d = 0; d < 3; ++d) { c.r_min[d] = p.r_min[d]; c.dr[d] = mg.drl(lvl)[d]; }
  c.nc = mg.bsl(lvl);
}

// src/m_build_tree.f90:18-174
void build_rectangle(Mg& mg, const int domain_size[3], int box_size, const double dx[3],
                     const double r_min[3], const bool periodic[3], int n_finer) {
  if (box_size % 2 != 0) fail("box_size should be even");
  for (int d = 0; d < 3; ++d)
    if (domain_size[d] % box_size != 0) fail("box_size does not divide domain_size");
  if (periodic[0] && periodic[1] && periodic[2]) mg.subtract_mean = true;

  int nx[3] = {domain_size[0], domain_size[1], domain_size[2]};
  int boxes_per_dim[NLVL][3];
  std::memset(boxes_per_dim, 0, sizeof(boxes_per_dim));
  mg.box_size = box_size;
  mg.bsl(1) = box_size;
  mg.first_normal_lvl = 1;
  for (int d = 0; d < 3; ++d) {
    mg.domain_size_lvl[1 - LVL_LO][d] = domain_size[d];
    mg.drl(1)[d] = dx[d];
    mg.r_min[d] = r_min[d];
    mg.periodic[d] = periodic[d];
    boxes_per_dim[1 - LVL_LO][d] = domain_size[d] / box_size;
  }

  int lvl;
  for (lvl = 1; lvl >= LVL_LO + 1; --lvl) {
    bool stop = false;
    for (int d = 0; d < 3; ++d)
      if ((nx[d] % 2 == 1) || nx[d] == mg.coarsest_grid[d] ||
          (mg.bsl(lvl) == mg.coarsest_grid[d] && mg.smoother_type == SM_GS)) stop = true;
    if (stop) break;
    bool all_even = true;
    for (int d = 0; d < 3; ++d) if ((nx[d] / mg.bsl(lvl)) % 2 != 0) all_even = false;
    if (all_even) {
      mg.bsl(lvl - 1) = mg.bsl(lvl);
      for (int d = 0; d < 3; ++d) boxes_per_dim[lvl - 1 - LVL_LO][d] = boxes_per_dim[lvl - LVL_LO][d] / 2;
      mg.first_normal_lvl = lvl - 1;
    } else {
      mg.bsl(lvl - 1) = mg.bsl(lvl) / 2;
      for (int d = 0; d < 3; ++d) boxes_per_dim[lvl - 1 - LVL_LO][d] = boxes_per_dim[lvl - LVL_LO][d];
    }
    for (int d = 0; d < 3; ++d) {
      mg.drl(lvl - 1)[d] = mg.drl(lvl)[d] * 2;
      nx[d] = nx[d] / 2;
      mg.domain_size_lvl[lvl - 1 - LVL_LO][d] = nx[d];
    }
  }
  mg.lowest_lvl = lvl;
  mg.highest_lvl = 1;

  for (int l = 2; l <= LVL_HI; ++l) {
    for (int d = 0; d < 3; ++d) {
      mg.drl(l)[d] = mg.drl(l - 1)[d] * 0.5;
      mg.domain_size_lvl[l - LVL_LO][d] = 2 * mg.domain_size_lvl[l - 1 - LVL_LO][d];
    }
    mg.bsl(l) = box_size;
  }

  long n = n_finer;
  for (int l = LVL_LO; l <= 1; ++l)
    n += (long)boxes_per_dim[l - LVL_LO][0] * boxes_per_dim[l - LVL_LO][1] * boxes_per_dim[l - LVL_LO][2];
  mg.boxes.assign((size_t)n + 1, Box());
  mg.n_boxes = 0;

  // lowest level, x fastest (KJI_DO_VEC)
  int* nb_ = boxes_per_dim[mg.lowest_lvl - LVL_LO];
  int nxx[3] = {nb_[0], nb_[1], nb_[2]};
  int periodic_offset[3] = {nxx[0] - 1, (nxx[1] - 1) * nxx[0], (nxx[2] - 1) * nxx[1] * nxx[0]};
  for (int k = 1; k <= nxx[2]; ++k) for (int j = 1; j <= nxx[1]; ++j) for (int i = 1; i <= nxx[0]; ++i) {
    mg.n_boxes += 1;
    int nn = mg.n_boxes;
    Box& b = mg.boxes[nn];
    b.rank = 0; b.id = nn; b.lvl = mg.lowest_lvl;
    b.ix[0] = i; b.ix[1] = j; b.ix[2] = k;
    for (int d = 0; d < 3; ++d) {
      b.r_min[d] = r_min[d] + (b.ix[d] - 1) * mg.bsl(mg.lowest_lvl) * mg.drl(mg.lowest_lvl)[d];
      b.dr[d] = mg.drl(mg.lowest_lvl)[d];
    }
    b.parent = NO_BOX;
    for (int c = 0; c < 8; ++c) b.children[c] = NO_BOX;
    b.neighbors[0] = nn - 1; b.neighbors[1] = nn + 1;
    b.neighbors[2] = nn - nxx[0]; b.neighbors[3] = nn + nxx[0];
    b.neighbors[4] = nn - nxx[0] * nxx[1]; b.neighbors[5] = nn + nxx[0] * nxx[1];
    int ijk[3] = {i, j, k};
    for (int d = 0; d < 3; ++d) {
      if (ijk[d] == 1) b.neighbors[2 * d] = periodic[d] ? nn + periodic_offset[d] : PHYSICAL_BOUNDARY;
      if (ijk[d] == nxx[d]) b.neighbors[2 * d + 1] = periodic[d] ? nn - periodic_offset[d] : PHYSICAL_BOUNDARY;
    }
    b.nc = mg.bsl(mg.lowest_lvl);
  }
  for (int l = 0; l < NLVL; ++l) mg.lvls[l] = Lvl();
  for (int nn = 1; nn <= mg.n_boxes; ++nn) mg.L(mg.lowest_lvl).ids.push_back(nn);

  for (int l = mg.lowest_lvl; l <= 0; ++l) {
    std::vector<int> ids = mg.L(l).ids;
    if (mg.bsl(l + 1) == mg.bsl(l)) {
      for (int id : ids) add_children(mg, id);
      set_leaves_parents(mg, l);
      set_next_level_ids(mg, l);
      set_neighbors_lvl(mg, l + 1);
    } else {
      for (int id : ids) add_single_child(mg, id, (int)ids.size());
      set_leaves_parents(mg, l);
      set_next_level_ids(mg, l);
    }
  }
  set_leaves_parents(mg, 1);
  for (int l = mg.lowest_lvl; l <= 1; ++l) mg.L(l).ref_bnds.clear();
  mg.tree_created = true;
}

// ===========================================================================
// Load balancing: src/m_load_balance.f90
// ===========================================================================

// src/m_load_balance.f90:197-221
int most_popular(const std::vector<int>& list, const std::vector<int>& work) {
  int best_count = 0, best_work = 0, res = -1;
  for (size_t i = 0; i < list.size(); ++i) {
    int cur = (int)std::count(list.begin(), list.end(), list[i]);
    int my_work = work[list[i]];
    if (cur > best_count || (cur == best_count && my_work < best_work)) {
      best_count = cur; best_work = my_work; res = list[i];
    }
  }
  return res;
}

// src/m_load_balance.f90:223-235
void update_lvl_info(Mg& mg) {
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    Lvl& L = mg.L(lvl);
    auto split = [&](const std::vector<int>& all, std::vector<std::vector<int>>& mine) {
      mine.assign(mg.n_cpu, std::vector<int>());
      for (int id : all) mine[mg.boxes[id].rank].push_back(id);
    };
    split(L.ids, L.my_ids); split(L.leaves, L.my_leaves);
    split(L.parents, L.my_parents); split(L.ref_bnds, L.my_ref_bnds);
  }
}

void set_coarse_rank(Mg& mg, int single_cpu_lvl, const std::vector<int>& my_work) {
  int coarse_rank = 0;
  if (single_cpu_lvl < mg.highest_lvl) {
    std::vector<int> ranks;
    for (int id : mg.L(single_cpu_lvl + 1).ids) ranks.push_back(mg.boxes[id].rank);
    coarse_rank = most_popular(ranks, my_work);
  }
  for (int lvl = mg.lowest_lvl; lvl <= single_cpu_lvl; ++lvl)
    for (int id : mg.L(lvl).ids) mg.boxes[id].rank = coarse_rank;
}

// src/m_load_balance.f90:22-63
void load_balance_simple(Mg& mg) {
  int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  for (int lvl = mg.lowest_lvl; lvl <= single_cpu_lvl; ++lvl)
    for (int id : mg.L(lvl).ids) mg.boxes[id].rank = 0;
  for (int lvl = single_cpu_lvl + 1; lvl <= mg.highest_lvl; ++lvl) {
    int work_left = (int)mg.L(lvl).ids.size(), my_work = 0, i_cpu = 0;
    for (int id : mg.L(lvl).ids) {
      if ((mg.n_cpu - i_cpu - 1) * my_work >= work_left) { i_cpu += 1; my_work = 0; }
      my_work += 1; work_left -= 1;
      mg.boxes[id].rank = i_cpu;
    }
  }
  update_lvl_info(mg);
}

// src/m_load_balance.f90:71-136
void load_balance(Mg& mg) {
  int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  std::vector<int> my_work(mg.n_cpu + 1, 0);
  for (int lvl = mg.highest_lvl; lvl >= single_cpu_lvl + 1; --lvl) {
    std::fill(my_work.begin(), my_work.end(), 0);
    for (int id : mg.L(lvl).parents) {
      std::vector<int> c_ranks;
      for (int c = 0; c < 8; ++c) c_ranks.push_back(mg.boxes[mg.boxes[id].children[c]].rank);
      int i_cpu = most_popular(c_ranks, my_work);
      mg.boxes[id].rank = i_cpu;
      my_work[i_cpu] += 1;
    }
    int work_left = (int)mg.L(lvl).leaves.size(), i_cpu = 0;
    for (int id : mg.L(lvl).leaves) {
      int rest = 0;
      for (int r = i_cpu + 1; r <= mg.n_cpu; ++r) rest += my_work[r];
      if ((mg.n_cpu - i_cpu - 1) * my_work[i_cpu] >= work_left + rest) i_cpu += 1;
      my_work[i_cpu] += 1; work_left -= 1;
      mg.boxes[id].rank = i_cpu;
    }
  }
  set_coarse_rank(mg, single_cpu_lvl, my_work);
  update_lvl_info(mg);
}

// src/m_load_balance.f90:140-193
void load_balance_parents(Mg& mg) {
  int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  std::vector<int> my_work(mg.n_cpu + 1, 0);
  for (int lvl = mg.highest_lvl - 1; lvl >= single_cpu_lvl + 1; --lvl) {
    std::fill(my_work.begin(), my_work.end(), 0);
    for (int id : mg.L(lvl).leaves) my_work[mg.boxes[id].rank] += 1;
    for (int id : mg.L(lvl).parents) {
      std::vector<int> c_ranks;
      for (int c = 0; c < 8; ++c) c_ranks.push_back(mg.boxes[mg.boxes[id].children[c]].rank);
      int i_cpu = most_popular(c_ranks, my_work);
      mg.boxes[id].rank = i_cpu;
      my_work[i_cpu] += 1;
    }
  }
  set_coarse_rank(mg, single_cpu_lvl, my_work);
  update_lvl_info(mg);
}

// src/m_allocate_storage.f90:51-99 (buffers are dynamic vectors here)
void allocate_storage(Mg& mg) {
  if (!mg.tree_created) fail("allocate_storage: tree is not yet created");
  if (mg.is_allocated) fail("allocate_storage: tree is already allocated");
  if (mg.L(mg.lowest_lvl).my_ids.empty()) update_lvl_info(mg);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    int nc = mg.bsl(lvl);
    for (int id : mg.L(lvl).ids) {
      Box& b = mg.boxes[id];
      b.nc = nc;
      b.cc.assign((size_t)(nc + 2) * (nc + 2) * (nc + 2) * mg.nvar(), 0.0);
    }
  }
  mg.buf.assign((size_t)mg.n_cpu * mg.n_cpu, Buf());
  mg.is_allocated = true;
}

// ===========================================================================
// Communication emulation: src/m_communication.f90:37-88
// ===========================================================================
void reset_bufs(Mg& mg) {
  for (Buf& b : mg.buf) { b.send.clear(); b.ix.clear(); b.i_recv = 0; }
}

// sort_sendbuf (:68-88): stable ascending rank by key (mrgrnk), then the
// transfer itself is the identity in shared memory: recv(dst<-src) == send(src->dst).
void sort_and_transfer_buffers(Mg& mg, int dsize) {
  for (Buf& b : mg.buf) {
    size_t n = b.ix.size();
    if (n == 0) continue;
    std::vector<int> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int c) { return b.ix[a] < b.ix[c]; });
    std::vector<double> cpy(b.send);
    std::vector<int> ixc(b.ix);
    for (size_t m = 0; m < n; ++m) {
      std::memcpy(&b.send[m * dsize], &cpy[(size_t)order[m] * dsize], sizeof(double) * dsize);
      b.ix[m] = ixc[order[m]];
    }
    b.i_recv = 0;
  }
}

// ===========================================================================
// Ghost cells: src/m_ghost_cells.f90
// ===========================================================================

// box_gc_for_neighbor :456-497  (gc(nc,nc), first tangential index fastest)
void box_gc_for_neighbor(const Box& box, int nb1, int nc, int iv, double* gc) {
  int n2 = nc + 2;
  int d = neighb_dim[nb1 - 1];
  int fix = neighb_low[nb1 - 1] ? 1 : nc;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    int i, j, k;
    if (d == 0) { i = fix; j = a; k = b; } else if (d == 1) { i = a; j = fix; k = b; } else { i = a; j = b; k = fix; }
    gc[(b - 1) * nc + (a - 1)] = box.cc[cix(n2, i, j, k, iv)];
  }
}

// box_get_gc :579-620 / box_set_gc :622-663
void box_get_gc(const Box& box, int nb1, int nc, int iv, double* gc) {
  int n2 = nc + 2;
  int d = neighb_dim[nb1 - 1];
  int fix = neighb_low[nb1 - 1] ? 0 : nc + 1;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    int i, j, k;
    if (d == 0) { i = fix; j = a; k = b; } else if (d == 1) { i = a; j = fix; k = b; } else { i = a; j = b; k = fix; }
    gc[(b - 1) * nc + (a - 1)] = box.cc[cix(n2, i, j, k, iv)];
  }
}
void box_set_gc(Box& box, int nb1, int nc, int iv, const double* gc) {
  int n2 = nc + 2;
  int d = neighb_dim[nb1 - 1];
  int fix = neighb_low[nb1 - 1] ? 0 : nc + 1;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    int i, j, k;
    if (d == 0) { i = fix; j = a; k = b; } else if (d == 1) { i = a; j = fix; k = b; } else { i = a; j = b; k = fix; }
    box.cc[cix(n2, i, j, k, iv)] = gc[(b - 1) * nc + (a - 1)];
  }
}

// box_gc_for_fine_neighbor :500-577
void box_gc_for_fine_neighbor(const Box& box, int nb1, const int di[3], int nc, int iv, double* gc) {
  int n2 = nc + 2, hnc = nc / 2, t2 = hnc + 2;
  std::vector<double> tmp((size_t)t2 * t2);
  int d = neighb_dim[nb1 - 1];
  int fix = neighb_low[nb1 - 1] ? 1 : nc;
  int ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;      // tangential dims (ascending)
  for (int b = 0; b <= hnc + 1; ++b) for (int a = 0; a <= hnc + 1; ++a) {
    int idx[3];
    idx[d] = fix; idx[ta] = di[ta] + a; idx[tb] = di[tb] + b;
    tmp[(size_t)b * t2 + a] = box.cc[cix(n2, idx[0], idx[1], idx[2], iv)];
  }
  auto T = [&](int a, int b) { return tmp[(size_t)b * t2 + a]; };
  auto G = [&](int a, int b) -> double& { return gc[(size_t)(b - 1) * nc + (a - 1)]; };
  for (int j = 1; j <= hnc; ++j) for (int i = 1; i <= hnc; ++i) {
    double g1 = 0.125 * (T(i + 1, j) - T(i - 1, j));
    double g2 = 0.125 * (T(i, j + 1) - T(i, j - 1));
    G(2 * i - 1, 2 * j - 1) = T(i, j) - g1 - g2;
    G(2 * i,     2 * j - 1) = T(i, j) + g1 - g2;
    G(2 * i - 1, 2 * j)     = T(i, j) - g1 + g2;
    G(2 * i,     2 * j)     = T(i, j) + g1 + g2;
  }
}

// bc_to_gc :665-766
void bc_to_gc(Mg& mg, int id, int nc, int iv, int nb1, int bc_type) {
  double c0, c1, c2;
  Box& box = mg.boxes[id];
  if (bc_type == BC_DIRICHLET) { c0 = 2; c1 = -1; c2 = 0; }
  else if (bc_type == BC_NEUMANN) {
    double dr = mg.drl(box.lvl)[neighb_dim[nb1 - 1]];
    c0 = dr * neighb_high_pm[nb1 - 1]; c1 = 1; c2 = 0;
  } else if (bc_type == BC_CONTINUOUS) { c0 = 0; c1 = 2; c2 = -1; }
  else fail("bc_to_gc: unknown boundary condition");
  int n2 = nc + 2, d = neighb_dim[nb1 - 1];
  bool low = neighb_low[nb1 - 1];
  int g = low ? 0 : nc + 1, i1 = low ? 1 : nc, i2 = low ? 2 : nc - 1;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    int p[3], q[3], r[3];
    int ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
    p[d] = g; q[d] = i1; r[d] = i2;
    p[ta] = q[ta] = r[ta] = a; p[tb] = q[tb] = r[tb] = b;
    double& gh = box.cc[cix(n2, p[0], p[1], p[2], iv)];
    gh = c0 * gh + c1 * box.cc[cix(n2, q[0], q[1], q[2], iv)] + c2 * box.cc[cix(n2, r[0], r[1], r[2], iv)];
  }
}

// sides_rb :769-861
void sides_rb(Box& box, int nc, int iv, int nb1, const double* gc) {
  int n2 = nc + 2, d = neighb_dim[nb1 - 1];
  int ix = neighb_low[nb1 - 1] ? 1 : nc, dix = neighb_low[nb1 - 1] ? 1 : -1;
  int ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    int p[3], q[3], r[3];
    p[d] = ix - dix; q[d] = ix; r[d] = ix + dix;
    p[ta] = q[ta] = r[ta] = a; p[tb] = q[tb] = r[tb] = b;
    box.cc[cix(n2, p[0], p[1], p[2], iv)] = 0.5 * gc[(size_t)(b - 1) * nc + (a - 1)]
        + 0.75 * box.cc[cix(n2, q[0], q[1], q[2], iv)]
        - 0.25 * box.cc[cix(n2, r[0], r[1], r[2], iv)];
  }
}

// buffer_for_nb :348-383, buffer_for_fine_nb :385-422 (sender side, rank `me`)
void buffer_ghost_cells(Mg& mg, int me, int id, int nc, int iv) {       // :177-198
  std::vector<double> gc((size_t)nc * nc);
  for (int nb = 1; nb <= 6; ++nb) {
    int nb_id = mg.boxes[id].neighbors[nb - 1];
    if (nb_id > NO_BOX) {
      int nb_rank = mg.boxes[nb_id].rank;
      if (nb_rank != me) {
        box_gc_for_neighbor(mg.boxes[id], nb, nc, iv, gc.data());
        Buf& b = mg.B(me, nb_rank);
        b.send.insert(b.send.end(), gc.begin(), gc.end());
        b.ix.push_back(6 * nb_id + neighb_rev[nb - 1]);
      }
    }
  }
}
void buffer_refinement_boundaries(Mg& mg, int me, int id, int nc, int iv) {   // :200-229
  std::vector<double> gc((size_t)nc * nc);
  for (int nb = 1; nb <= 6; ++nb) {
    int nb_id = mg.boxes[id].neighbors[nb - 1];
    if (nb_id > NO_BOX && has_children(mg.boxes[nb_id])) {
      for (int n = 0; n < 4; ++n) {
        int c_id = mg.boxes[nb_id].children[child_adj_nb[neighb_rev[nb - 1] - 1][n] - 1];
        int c_rank = mg.boxes[c_id].rank;
        if (c_rank != me) {
          int off[3]; get_child_offset(mg, c_id, off);
          box_gc_for_fine_neighbor(mg.boxes[id], nb, off, nc, iv, gc.data());
          Buf& b = mg.B(me, c_rank);
          b.send.insert(b.send.end(), gc.begin(), gc.end());
          b.ix.push_back(6 * c_id + neighb_rev[nb - 1]);
        }
      }
    }
  }
}

// set_ghost_cells :232-285 with fill_refinement_bnd :287-328, copy_from_nb :330-346,
// fill_buffered_nb :424-454 (receiver side, rank `me`)
void set_ghost_cells(Mg& mg, int me, int id, int nc, int iv) {
  std::vector<double> gc((size_t)nc * nc);
  int dsize = nc * nc;
  Box& box = mg.boxes[id];
  for (int nb = 1; nb <= 6; ++nb) {
    int nb_id = box.neighbors[nb - 1];
    if (nb_id > NO_BOX) {
      int nb_rank = mg.boxes[nb_id].rank;
      if (nb_rank != me) {
        Buf& b = mg.B(nb_rank, me);
        box_set_gc(box, nb, nc, iv, &b.send[b.i_recv]);
        b.i_recv += dsize;
      } else {
        box_gc_for_neighbor(mg.boxes[nb_id], neighb_rev[nb - 1], nc, iv, gc.data());
        box_set_gc(box, nb, nc, iv, gc.data());
      }
    } else if (nb_id == NO_BOX) {
      int p_id = box.parent;
      int p_nb_id = mg.boxes[p_id].neighbors[nb - 1];
      int p_nb_rank = mg.boxes[p_nb_id].rank;
      if (p_nb_rank != me) {
        Buf& b = mg.B(p_nb_rank, me);
        std::memcpy(gc.data(), &b.send[b.i_recv], sizeof(double) * dsize);
        b.i_recv += dsize;
      } else {
        int off[3]; get_child_offset(mg, id, off);
        box_gc_for_fine_neighbor(mg.boxes[p_nb_id], neighb_rev[nb - 1], off, nc, iv, gc.data());
      }
      sides_rb(box, nc, iv, nb, gc.data());
    } else {
      int bc_type;
      if (mg.phi_bc_data_stored && iv == IPHI) {
        box_get_gc(box, nb, nc, IRHS, gc.data());
        bc_type = nb_id;
      } else if (mg.bc[nb - 1][iv].boundary_cond) {
        const Bc& c = mg.bc[nb - 1][iv];
        c.boundary_cond(box.r_min, box.dr, box.lvl, nc, iv + 1, nb, &bc_type, gc.data(), c.user);
      } else {
        bc_type = mg.bc[nb - 1][iv].bc_type;
        std::fill(gc.begin(), gc.end(), mg.bc[nb - 1][iv].bc_value);
      }
      box_set_gc(box, nb, nc, iv, gc.data());
      bc_to_gc(mg, id, nc, iv, nb, bc_type);
    }
  }
}

// mg_fill_ghost_cells_lvl :131-175
void fill_ghost_cells_lvl(Mg& mg, int lvl, int iv) {
  if (lvl < mg.lowest_lvl) fail("fill_ghost_cells_lvl: lvl < lowest_lvl");
  if (lvl > mg.highest_lvl) fail("fill_ghost_cells_lvl: lvl > highest_lvl");
  int nc = mg.bsl(lvl);
  if (lvl >= mg.first_normal_lvl && mg.n_cpu > 1) {
    reset_bufs(mg);
    #pragma omp parallel for schedule(static, 1)
    for (int me = 0; me < mg.n_cpu; ++me) {
      for (int id : mg.L(lvl).my_ids[me]) buffer_ghost_cells(mg, me, id, nc, iv);
      if (lvl > 1)
        for (int id : mg.L(lvl - 1).my_ref_bnds[me]) buffer_refinement_boundaries(mg, me, id, nc, iv);
    }
    sort_and_transfer_buffers(mg, nc * nc);
  }
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl).my_ids[me]) set_ghost_cells(mg, me, id, nc, iv);
}

// mg_phi_bc_store :66-117
void phi_bc_store(Mg& mg) {
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    int nc = mg.bsl(lvl);
    std::vector<double> bc((size_t)nc * nc);
    for (int id : mg.L(lvl).ids) {
      Box& box = mg.boxes[id];
      for (int nb = 1; nb <= 6; ++nb) {
        if (box.neighbors[nb - 1] < NO_BOX) {
          int bc_type;
          const Bc& c = mg.bc[nb - 1][IPHI];
          if (c.boundary_cond) c.boundary_cond(box.r_min, box.dr, box.lvl, nc, IPHI + 1, nb, &bc_type, bc.data(), c.user);
          else { bc_type = c.bc_type; std::fill(bc.begin(), bc.end(), c.bc_value); }
          box.neighbors[nb - 1] = bc_type;
          box_set_gc(box, nb, nc, IRHS, bc.data());
        }
      }
    }
  }
  mg.phi_bc_data_stored = true;
}

// ===========================================================================
// Operators and smoothers (3-D Cartesian)
// ===========================================================================
struct Geo { double idr2[3]; };
inline Geo geo(Mg& mg, int lvl) {
  Geo g; for (int d = 0; d < 3; ++d) { double dr = mg.drl(lvl)[d]; g.idr2[d] = 1 / (dr * dr); } return g;
}

// box_lpl src/m_laplacian.f90:155-195; box_helmh src/m_helmholtz.f90:111-154
void box_lpl(Mg& mg, int id, int nc, int i_out, bool helm) {
  Box& b = mg.boxes[id]; int n2 = nc + 2; Geo g = geo(mg, b.lvl);
  double* cc = b.cc.data(); const int n = IPHI; double lambda = mg.helmholtz_lambda;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
    double c0 = cc[cix(n2, i, j, k, n)];
    double v = g.idr2[0] * (cc[cix(n2, i - 1, j, k, n)] + cc[cix(n2, i + 1, j, k, n)] - 2 * c0)
             + g.idr2[1] * (cc[cix(n2, i, j - 1, k, n)] + cc[cix(n2, i, j + 1, k, n)] - 2 * c0)
             + g.idr2[2] * (cc[cix(n2, i, j, k - 1, n)] + cc[cix(n2, i, j, k + 1, n)] - 2 * c0);
    if (helm) v = v - lambda * c0;
    cc[cix(n2, i, j, k, i_out)] = v;
  }
}

// box_gs_lpl src/m_laplacian.f90:52-114; box_gs_helmh src/m_helmholtz.f90:49-108
void box_gs_lpl(Mg& mg, int id, int nc, int cntr, bool helm) {
  Box& b = mg.boxes[id]; int n2 = nc + 2; Geo g = geo(mg, b.lvl);
  double s = g.idr2[0] + g.idr2[1] + g.idr2[2];
  double fac = helm ? 1.0 / (2 * s + mg.helmholtz_lambda) : 0.5 / s;
  bool rb = (mg.smoother_type == SM_GSRB);
  int di = rb ? 2 : 1;
  double* cc = b.cc.data(); const int n = IPHI;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) {
    int i0 = rb ? 2 - ((cntr ^ (k + j)) & 1) : 1;
    for (int i = i0; i <= nc; i += di) {
      cc[cix(n2, i, j, k, n)] = fac * (
          g.idr2[0] * (cc[cix(n2, i + 1, j, k, n)] + cc[cix(n2, i - 1, j, k, n)]) +
          g.idr2[1] * (cc[cix(n2, i, j + 1, k, n)] + cc[cix(n2, i, j - 1, k, n)]) +
          g.idr2[2] * (cc[cix(n2, i, j, k + 1, n)] + cc[cix(n2, i, j, k - 1, n)]) -
          cc[cix(n2, i, j, k, IRHS)]);
    }
  }
}

// gather the 6 neighbours in reference face order (-x,+x,-y,+y,-z,+z)
inline void gather6(const double* cc, int n2, int i, int j, int k, int v, double* u) {
  u[0] = cc[cix(n2, i - 1, j, k, v)]; u[1] = cc[cix(n2, i + 1, j, k, v)];
  u[2] = cc[cix(n2, i, j - 1, k, v)]; u[3] = cc[cix(n2, i, j + 1, k, v)];
  u[4] = cc[cix(n2, i, j, k - 1, v)]; u[5] = cc[cix(n2, i, j, k + 1, v)];
}

// box_vlpl src/m_vlaplacian.f90:131-189; box_vhelmh src/m_vhelmholtz.f90:144-205;
// box_ahelmh src/m_ahelmholtz.f90:165-237 (aniso: eps vars 5,6,7 per direction)
void box_vlpl(Mg& mg, int id, int nc, int i_out, bool helm, bool aniso) {
  Box& b = mg.boxes[id]; int n2 = nc + 2; Geo g = geo(mg, b.lvl);
  double idr2[6] = {g.idr2[0], g.idr2[0], g.idr2[1], g.idr2[1], g.idr2[2], g.idr2[2]};
  double* cc = b.cc.data(); double lambda = mg.helmholtz_lambda;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
    double u0 = cc[cix(n2, i, j, k, IPHI)], u[6], a[6], a0[6];
    gather6(cc, n2, i, j, k, IPHI, u);
    if (!aniso) {
      gather6(cc, n2, i, j, k, IVEPS, a);
      for (int f = 0; f < 6; ++f) a0[f] = cc[cix(n2, i, j, k, IVEPS)];
    } else {
      double t[6];
      for (int d = 0; d < 3; ++d) {
        gather6(cc, n2, i, j, k, IVEPS + d, t);
        a[2 * d] = t[2 * d]; a[2 * d + 1] = t[2 * d + 1];
        a0[2 * d] = a0[2 * d + 1] = cc[cix(n2, i, j, k, IVEPS + d)];
      }
    }
    double s = 0.0;
    for (int f = 0; f < 6; ++f) s = s + 2 * idr2[f] * a0[f] * a[f] / (a0[f] + a[f]) * (u[f] - u0);
    if (helm) s = s - lambda * u0;
    cc[cix(n2, i, j, k, i_out)] = s;
  }
}

// box_gs_vlpl src/m_vlaplacian.f90:51-128; box_gs_vhelmh src/m_vhelmholtz.f90:61-141;
// box_gs_ahelmh src/m_ahelmholtz.f90:69-162.  NOTE: the reference's 3-D
// anisotropic smoother writes a0(4:5)=eps3 leaving a0(6) uninitialised; the evidently
// intended a0(5:6)=eps3 is implemented here (matches box_ahelmh).
void box_gs_vlpl(Mg& mg, int id, int nc, int cntr, bool helm, bool aniso) {
  Box& b = mg.boxes[id]; int n2 = nc + 2; Geo g = geo(mg, b.lvl);
  double idr2[6] = {g.idr2[0], g.idr2[0], g.idr2[1], g.idr2[1], g.idr2[2], g.idr2[2]};
  bool rb = (mg.smoother_type == SM_GSRB);
  int di = rb ? 2 : 1;
  double* cc = b.cc.data(); double lambda = mg.helmholtz_lambda;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) {
    int i0 = rb ? 2 - ((cntr ^ (k + j)) & 1) : 1;
    for (int i = i0; i <= nc; i += di) {
      double u[6], a[6], a0[6], c[6];
      gather6(cc, n2, i, j, k, IPHI, u);
      if (!aniso) {
        gather6(cc, n2, i, j, k, IVEPS, a);
        for (int f = 0; f < 6; ++f) a0[f] = cc[cix(n2, i, j, k, IVEPS)];
      } else {
        double t[6];
        for (int d = 0; d < 3; ++d) {
          gather6(cc, n2, i, j, k, IVEPS + d, t);
          a[2 * d] = t[2 * d]; a[2 * d + 1] = t[2 * d + 1];
          a0[2 * d] = a0[2 * d + 1] = cc[cix(n2, i, j, k, IVEPS + d)];
        }
      }
      for (int f = 0; f < 6; ++f) c[f] = 2 * a0[f] * a[f] / (a0[f] + a[f]) * idr2[f];
      double scu = 0.0, sc = 0.0;
      for (int f = 0; f < 6; ++f) scu = scu + c[f] * u[f];
      for (int f = 0; f < 6; ++f) sc = sc + c[f];
      double den = helm ? (sc + lambda) : sc;
      cc[cix(n2, i, j, k, IPHI)] = (scu - cc[cix(n2, i, j, k, IRHS)]) / den;
    }
  }
}

void box_op(Mg& mg, int id, int nc, int i_out) {
  switch (mg.operator_type) {
    case OP_LAPLACIAN: box_lpl(mg, id, nc, i_out, false); break;
    case OP_HELMHOLTZ: box_lpl(mg, id, nc, i_out, true); break;
    case OP_VLAPLACIAN: box_vlpl(mg, id, nc, i_out, false, false); break;
    case OP_VHELMHOLTZ: box_vlpl(mg, id, nc, i_out, true, false); break;
    case OP_AHELMHOLTZ: box_vlpl(mg, id, nc, i_out, true, true); break;
    default: fail("mg_set_methods: unknown operator");
  }
}
void box_smoother(Mg& mg, int id, int nc, int cntr) {
  switch (mg.operator_type) {
    case OP_LAPLACIAN: box_gs_lpl(mg, id, nc, cntr, false); break;
    case OP_HELMHOLTZ: box_gs_lpl(mg, id, nc, cntr, true); break;
    case OP_VLAPLACIAN: box_gs_vlpl(mg, id, nc, cntr, false, false); break;
    case OP_VHELMHOLTZ: box_gs_vlpl(mg, id, nc, cntr, true, false); break;
    case OP_AHELMHOLTZ: box_gs_vlpl(mg, id, nc, cntr, true, true); break;
    default: fail("mg_set_methods: unknown operator");
  }
}

// mg_set_methods src/m_multigrid.f90:27-60 + *_set_methods of the operator modules
void set_methods(Mg& mg) {
  bool all_periodic = mg.periodic[0] && mg.periodic[1] && mg.periodic[2];
  auto neumann_eps = [&](int iv) { for (int nb = 0; nb < 6; ++nb) { mg.bc[nb][iv].bc_type = BC_NEUMANN; mg.bc[nb][iv].bc_value = 0.0; } };
  auto need_extra = [&](int n) {
    if (mg.n_extra_vars == 0 && mg.is_allocated) fail("set_methods: mg%n_extra_vars == 0");
    mg.n_extra_vars = std::max(n, mg.n_extra_vars);
  };
  switch (mg.operator_type) {
    case OP_LAPLACIAN: if (all_periodic) mg.subtract_mean = true; break;      // m_laplacian.f90:16-20
    case OP_VLAPLACIAN: need_extra(1); mg.subtract_mean = false; neumann_eps(IVEPS); break;
    case OP_HELMHOLTZ: mg.subtract_mean = false; break;
    case OP_VHELMHOLTZ: need_extra(1); mg.subtract_mean = false; neumann_eps(IVEPS); break;
    case OP_AHELMHOLTZ: need_extra(3); neumann_eps(IVEPS); neumann_eps(IVEPS + 1); neumann_eps(IVEPS + 2); break;
    default: fail("mg_set_methods: unknown operator");
  }
  if (mg.smoother_type != SM_GS && mg.smoother_type != SM_GSRB) fail("set_methods: unsupported smoother type");
  mg.n_smoother_substeps = (mg.smoother_type == SM_GSRB) ? 2 : 1;
  mg.methods_set = true;
}

// ===========================================================================
// Restriction: src/m_restrict.f90
// ===========================================================================
inline double restrict_cell(const double* cc, int n2, int i, int j, int k, int iv) {
  // 0.125 * sum(cc(2i-1:2i, 2j-1:2j, 2k-1:2k)), Fortran array-element order
  double s = 0.0;
  for (int kk = 2 * k - 1; kk <= 2 * k; ++kk) for (int jj = 2 * j - 1; jj <= 2 * j; ++jj)
    for (int ii = 2 * i - 1; ii <= 2 * i; ++ii) s = s + cc[cix(n2, ii, jj, kk, iv)];
  return 0.125 * s;
}

// restrict_set_buffer :116-163
void restrict_set_buffer(Mg& mg, int me, int id, int iv) {
  int hnc = mg.box_size / 2;
  int p_id = mg.boxes[id].parent, p_rank = mg.boxes[p_id].rank;
  if (p_rank == me) return;
  const Box& b = mg.boxes[id]; int n2 = b.nc + 2;
  Buf& buf = mg.B(me, p_rank);
  for (int k = 1; k <= hnc; ++k) for (int j = 1; j <= hnc; ++j) for (int i = 1; i <= hnc; ++i)
    buf.send.push_back(restrict_cell(b.cc.data(), n2, i, j, k, iv));
  buf.ix.push_back(8 * p_id + ix_to_ichild(b.ix));
}

// restrict_onto :165-214
void restrict_onto(Mg& mg, int me, int id, int nc, int iv) {
  int hnc = nc / 2, dsize = hnc * hnc * hnc;
  Box& p = mg.boxes[id]; int pn2 = p.nc + 2;
  for (int i_c = 0; i_c < 8; ++i_c) {
    int c_id = p.children[i_c];
    if (c_id == NO_BOX) continue;
    int c_rank = mg.boxes[c_id].rank;
    int dix[3]; get_child_offset(mg, c_id, dix);
    if (c_rank == me) {
      const Box& c = mg.boxes[c_id]; int n2 = c.nc + 2;
      for (int j = 1; j <= hnc; ++j) for (int i = 1; i <= hnc; ++i) for (int k = 1; k <= hnc; ++k)
        p.cc[cix(pn2, dix[0] + i, dix[1] + j, dix[2] + k, iv)] = restrict_cell(c.cc.data(), n2, i, j, k, iv);
    } else {
      Buf& buf = mg.B(c_rank, me);
      const double* r = &buf.send[buf.i_recv];
      size_t m = 0;
      for (int k = 1; k <= hnc; ++k) for (int j = 1; j <= hnc; ++j) for (int i = 1; i <= hnc; ++i)
        p.cc[cix(pn2, dix[0] + i, dix[1] + j, dix[2] + k, iv)] = r[m++];
      buf.i_recv += dsize;
    }
  }
}

// mg_restrict_lvl :83-114
void restrict_lvl(Mg& mg, int iv, int lvl) {
  if (lvl <= mg.lowest_lvl) fail("cannot restrict lvl <= lowest_lvl");
  int nc = mg.bsl(lvl);
  if (lvl >= mg.first_normal_lvl && mg.n_cpu > 1) {
    reset_bufs(mg);
    #pragma omp parallel for schedule(static, 1)
    for (int me = 0; me < mg.n_cpu; ++me)
      for (int id : mg.L(lvl).my_ids[me]) restrict_set_buffer(mg, me, id, iv);
    int h = nc / 2;
    sort_and_transfer_buffers(mg, h * h * h);
  }
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl - 1).my_parents[me]) restrict_onto(mg, me, id, nc, iv);
}

// ===========================================================================
// Prolongation: src/m_prolong.f90
// ===========================================================================

// mg_prolong_sparse :159-240
void prolong_sparse(const Mg& mg, int p_id, const int dix[3], int nc, int iv, double* fine) {
  const Box& p = mg.boxes[p_id]; int n2 = p.nc + 2, hnc = nc / 2;
  const double* crs = p.cc.data();
  auto F = [&](int i, int j, int k) -> double& { return fine[((size_t)(k - 1) * nc + (j - 1)) * nc + (i - 1)]; };
  for (int k = 1; k <= hnc; ++k) { int kc = k + dix[2];
    for (int j = 1; j <= hnc; ++j) { int jc = j + dix[1];
      for (int i = 1; i <= hnc; ++i) { int ic = i + dix[0];
        double f0  = 0.25 * crs[cix(n2, ic, jc, kc, iv)];
        double flx = 0.25 * crs[cix(n2, ic - 1, jc, kc, iv)];
        double fhx = 0.25 * crs[cix(n2, ic + 1, jc, kc, iv)];
        double fly = 0.25 * crs[cix(n2, ic, jc - 1, kc, iv)];
        double fhy = 0.25 * crs[cix(n2, ic, jc + 1, kc, iv)];
        double flz = 0.25 * crs[cix(n2, ic, jc, kc - 1, iv)];
        double fhz = 0.25 * crs[cix(n2, ic, jc, kc + 1, iv)];
        F(2*i-1, 2*j-1, 2*k-1) = f0 + flx + fly + flz;
        F(2*i,   2*j-1, 2*k-1) = f0 + fhx + fly + flz;
        F(2*i-1, 2*j,   2*k-1) = f0 + flx + fhy + flz;
        F(2*i,   2*j,   2*k-1) = f0 + fhx + fhy + flz;
        F(2*i-1, 2*j-1, 2*k)   = f0 + flx + fly + fhz;
        F(2*i,   2*j-1, 2*k)   = f0 + fhx + fly + fhz;
        F(2*i-1, 2*j,   2*k)   = f0 + flx + fhy + fhz;
        F(2*i,   2*j,   2*k)   = f0 + fhx + fhy + fhz;
      } } }
}

// prolong_set_buffer :91-121
void prolong_set_buffer(Mg& mg, int me, int id, int nc, int iv) {
  std::vector<double> tmp((size_t)nc * nc * nc);
  for (int i_c = 0; i_c < 8; ++i_c) {
    int c_id = mg.boxes[id].children[i_c];
    if (c_id > NO_BOX) {
      int c_rank = mg.boxes[c_id].rank;
      if (c_rank != me) {
        int dix[3]; get_child_offset(mg, c_id, dix);
        prolong_sparse(mg, id, dix, nc, iv, tmp.data());
        Buf& b = mg.B(me, c_rank);
        b.send.insert(b.send.end(), tmp.begin(), tmp.end());
        b.ix.push_back(c_id);
      }
    }
  }
}

// prolong_onto :124-156
void prolong_onto(Mg& mg, int me, int id, int nc, int iv, int iv_to, bool add) {
  std::vector<double> tmp((size_t)nc * nc * nc);
  Box& b = mg.boxes[id];
  int p_id = b.parent, p_rank = mg.boxes[p_id].rank;
  if (p_rank == me) {
    int dix[3]; get_child_offset(mg, id, dix);
    prolong_sparse(mg, p_id, dix, nc, iv, tmp.data());
  } else {
    Buf& buf = mg.B(p_rank, me);
    std::memcpy(tmp.data(), &buf.send[buf.i_recv], sizeof(double) * tmp.size());
    buf.i_recv += tmp.size();
  }
  int n2 = nc + 2; size_t m = 0;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i, ++m) {
    double& t = b.cc[cix(n2, i, j, k, iv_to)];
    t = add ? t + tmp[m] : tmp[m];
  }
}

// mg_prolong :51-85
void prolong(Mg& mg, int lvl, int iv, int iv_to, bool add) {
  if (lvl == mg.highest_lvl) fail("cannot prolong highest level");
  if (lvl < mg.lowest_lvl) fail("cannot prolong below lowest level");
  if (lvl >= mg.first_normal_lvl - 1 && mg.n_cpu > 1) {
    reset_bufs(mg);
    #pragma omp parallel for schedule(static, 1)
    for (int me = 0; me < mg.n_cpu; ++me)
      for (int id : mg.L(lvl).my_ids[me]) prolong_set_buffer(mg, me, id, mg.box_size, iv);
    sort_and_transfer_buffers(mg, mg.box_size * mg.box_size * mg.box_size);
  }
  int nc = mg.bsl(lvl + 1);
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl + 1).my_ids[me]) prolong_onto(mg, me, id, nc, iv, iv_to, add);
}

// ===========================================================================
// Multigrid cycle: src/m_multigrid.f90
// ===========================================================================

// residual_box :426-436
void residual_box(Mg& mg, int id, int nc) {
  box_op(mg, id, nc, IRES);
  Box& b = mg.boxes[id]; int n2 = nc + 2;
  for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
    double& r = b.cc[cix(n2, i, j, k, IRES)];
    r = b.cc[cix(n2, i, j, k, IRHS)] - r;
  }
}

// max_residual_lvl :296-311 (per-rank maxima; max is order independent)
double max_residual_lvl(Mg& mg, int lvl) {
  int nc = mg.bsl(lvl);
  std::vector<double> rmax(mg.n_cpu, 0.0);
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me) {
    double m = 0.0;
    for (int id : mg.L(lvl).my_ids[me]) {
      residual_box(mg, id, nc);
      const Box& b = mg.boxes[id]; int n2 = nc + 2;
      for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i)
        m = std::max(m, std::fabs(b.cc[cix(n2, i, j, k, IRES)]));
    }
    rmax[me] = m;
  }
  return *std::max_element(rmax.begin(), rmax.end());
}

// smooth_boxes :404-424
void smooth_boxes(Mg& mg, int lvl, int n_cycle) {
  int nc = mg.bsl(lvl);
  for (int n = 1; n <= n_cycle * mg.n_smoother_substeps; ++n) {
    tstart(mg, T_SMOOTH);
    #pragma omp parallel for schedule(static, 1)
    for (int me = 0; me < mg.n_cpu; ++me)
      for (int id : mg.L(lvl).my_ids[me]) box_smoother(mg, id, nc, n);
    tend(mg, T_SMOOTH);
    tstart(mg, T_SMOOTH_GC);
    fill_ghost_cells_lvl(mg, lvl, IPHI);
    tend(mg, T_SMOOTH_GC);
  }
}

// update_coarse :347-384
void update_coarse(Mg& mg, int lvl) {
  int nc = mg.bsl(lvl), nc_c = mg.bsl(lvl - 1);
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl).my_ids[me]) residual_box(mg, id, nc);
  restrict_lvl(mg, IPHI, lvl);
  restrict_lvl(mg, IRES, lvl);
  fill_ghost_cells_lvl(mg, lvl - 1, IPHI);
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl - 1).my_parents[me]) {
      box_op(mg, id, nc_c, IRHS);
      Box& b = mg.boxes[id]; int n2 = nc_c + 2;
      for (int k = 1; k <= nc_c; ++k) for (int j = 1; j <= nc_c; ++j) for (int i = 1; i <= nc_c; ++i)
        b.cc[cix(n2, i, j, k, IRHS)] = b.cc[cix(n2, i, j, k, IRHS)] + b.cc[cix(n2, i, j, k, IRES)];
      size_t n3 = (size_t)n2 * n2 * n2;
      std::memcpy(&b.cc[n3 * IOLD], &b.cc[n3 * IPHI], sizeof(double) * n3);
    }
}

// correct_children :387-402
void correct_children(Mg& mg, int lvl) {
  #pragma omp parallel for schedule(static, 1)
  for (int me = 0; me < mg.n_cpu; ++me)
    for (int id : mg.L(lvl).my_parents[me]) {
      Box& b = mg.boxes[id]; int n2 = b.nc + 2; size_t n3 = (size_t)n2 * n2 * n2;
      for (size_t m = 0; m < n3; ++m) b.cc[n3 * IRES + m] = b.cc[n3 * IPHI + m] - b.cc[n3 * IOLD + m];
    }
  prolong(mg, lvl, IRES, IPHI, true);
}

// get_sum :278-294 and subtract_mean :245-276.  The cross-rank sum is the
// reference's MPI_Allreduce(SUM): rank partial sums added in rank order.
void subtract_mean(Mg& mg, int iv, bool include_ghostcells) {
  double total = 0.0;
  for (int me = 0; me < mg.n_cpu; ++me) {
    double s = 0.0;
    for (int lvl = 1; lvl <= mg.highest_lvl; ++lvl) {
      int nc = mg.bsl(lvl), n2 = nc + 2;
      double w = mg.drl(lvl)[0] * mg.drl(lvl)[1] * mg.drl(lvl)[2];
      for (int id : mg.L(lvl).my_leaves[me]) {
        const Box& b = mg.boxes[id];
        double bs = 0.0;
        for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i)
          bs = bs + b.cc[cix(n2, i, j, k, iv)];
        s = s + w * bs;
      }
    }
    total = total + s;
  }
  int nc = mg.box_size;
  double volume = (double)(nc * nc * nc) * (mg.drl(1)[0] * mg.drl(1)[1] * mg.drl(1)[2]) * (double)mg.L(1).ids.size();
  double mean = total / volume;
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    int ncl = mg.bsl(lvl), n2 = ncl + 2; size_t n3 = (size_t)n2 * n2 * n2;
    for (int id : mg.L(lvl).ids) {
      Box& b = mg.boxes[id];
      if (include_ghostcells) { for (size_t m = 0; m < n3; ++m) b.cc[n3 * iv + m] -= mean; }
      else for (int k = 1; k <= ncl; ++k) for (int j = 1; j <= ncl; ++j) for (int i = 1; i <= ncl; ++i)
        b.cc[cix(n2, i, j, k, iv)] -= mean;
    }
  }
}

// mg_fas_vcycle :150-243
void fas_vcycle(Mg& mg, bool has_highest, int highest_lvl, bool want_res, double* max_res, bool standalone) {
  if (!mg.methods_set) set_methods(mg);
  if (standalone) tstart(mg, T_VCYCLE);
  if (mg.subtract_mean && !has_highest) subtract_mean(mg, IRHS, false);
  int min_lvl = mg.lowest_lvl, max_lvl = has_highest ? highest_lvl : mg.highest_lvl;
  if (standalone) fill_ghost_cells_lvl(mg, max_lvl, IPHI);

  for (int lvl = max_lvl; lvl >= min_lvl + 1; --lvl) {
    smooth_boxes(mg, lvl, mg.n_cycle_down);
    tstart(mg, T_UPDATE_COARSE);
    update_coarse(mg, lvl);
    tend(mg, T_UPDATE_COARSE);
  }

  tstart(mg, T_COARSE);
  {
    const std::vector<int>& ids = mg.L(min_lvl).ids;
    for (int id : ids) if (mg.boxes[id].rank != mg.boxes[ids[0]].rank)
      fail("Multiple CPUs for coarse grid (not implemented yet)");
  }
  double init_res = max_residual_lvl(mg, min_lvl);
  int it = 0;
  for (int i = 1; i <= mg.max_coarse_cycles; ++i) {
    smooth_boxes(mg, min_lvl, mg.n_cycle_up + mg.n_cycle_down);
    double res = max_residual_lvl(mg, min_lvl);
    it = i;
    if (res < mg.residual_coarse_rel * init_res || res < mg.residual_coarse_abs) break;
  }
  mg.last_coarse_iters = it;
  tend(mg, T_COARSE);

  for (int lvl = min_lvl + 1; lvl <= max_lvl; ++lvl) {
    tstart(mg, T_CORRECT);
    correct_children(mg, lvl - 1);
    fill_ghost_cells_lvl(mg, lvl, IPHI);
    tend(mg, T_CORRECT);
    smooth_boxes(mg, lvl, mg.n_cycle_up);
  }

  if (want_res) {
    double r = 0.0;
    for (int lvl = min_lvl; lvl <= max_lvl; ++lvl) r = std::max(max_residual_lvl(mg, lvl), r);
    *max_res = r;
  }
  if (mg.subtract_mean) subtract_mean(mg, IPHI, true);
  if (standalone) tend(mg, T_VCYCLE);
}

// mg_fas_fmg :84-147
void fas_fmg(Mg& mg, bool have_guess, bool want_res, double* max_res) {
  if (!mg.methods_set) set_methods(mg);
  tstart(mg, T_FMG);
  if (!have_guess) {
    for (int lvl = mg.highest_lvl; lvl >= mg.lowest_lvl; --lvl)
      for (int id : mg.L(lvl).ids) {
        Box& b = mg.boxes[id]; int n2 = b.nc + 2; size_t n3 = (size_t)n2 * n2 * n2;
        std::fill(b.cc.begin() + n3 * IPHI, b.cc.begin() + n3 * (IPHI + 1), 0.0);
      }
  }
  fill_ghost_cells_lvl(mg, mg.highest_lvl, IPHI);
  for (int lvl = mg.highest_lvl; lvl >= mg.lowest_lvl + 1; --lvl) {
    tstart(mg, T_UPDATE_COARSE);
    update_coarse(mg, lvl);
    tend(mg, T_UPDATE_COARSE);
  }
  if (mg.subtract_mean) subtract_mean(mg, IRHS, false);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    for (int id : mg.L(lvl).ids) {
      Box& b = mg.boxes[id]; int n2 = b.nc + 2; size_t n3 = (size_t)n2 * n2 * n2;
      std::memcpy(&b.cc[n3 * IOLD], &b.cc[n3 * IPHI], sizeof(double) * n3);
    }
    if (lvl > mg.lowest_lvl) {
      tstart(mg, T_CORRECT);
      correct_children(mg, lvl - 1);
      tend(mg, T_CORRECT);
      fill_ghost_cells_lvl(mg, lvl, IPHI);
    }
    if (lvl == mg.highest_lvl) fas_vcycle(mg, true, lvl, want_res, max_res, false);
    else fas_vcycle(mg, true, lvl, false, nullptr, false);
  }
  tend(mg, T_FMG);
}

// mg_apply_op :439-456
void apply_op(Mg& mg, int i_out) {
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl)
    for (int id : mg.L(lvl).ids) box_op(mg, id, mg.bsl(lvl), i_out);
}

// ===========================================================================
// Implicit diffusion: src/m_diffusion.f90 (the production caller of the cycle)
// ===========================================================================
// set_rhs :159-173 -- leaves of levels 1..highest only, interior cells
void diffusion_set_rhs(Mg& mg, double f1, double f2) {
  for (int lvl = 1; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.bsl(lvl), n2 = nc + 2;
    for (int id : mg.L(lvl).leaves) {
      double* cc = mg.boxes[id].cc.data();
      for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
        const double a = f1 * cc[cix(n2, i, j, k, IPHI)];
        const double b = f2 * cc[cix(n2, i, j, k, IRHS)];
        cc[cix(n2, i, j, k, IRHS)] = a + b;
      }
    }
  }
}

// diffusion_solve :19-58 (op = helmholtz, coefficient folded into lambda), diffusion_solve_vcoeff :64-106,
// diffusion_solve_acoeff :113-157 (lambda = 1/dt or 2/dt).  Returns the number of V-cycles added after the FMG cycle.
int diffusion_solve(Mg& mg, int op, double dt, double diffusion_coeff, int order, double max_res, double* res_out) {
  const int max_its = 10;
  if (op != OP_HELMHOLTZ && op != OP_VHELMHOLTZ && op != OP_AHELMHOLTZ) fail("diffusion_solve: operator must be a Helmholtz variant");
  mg.operator_type = op;
  set_methods(mg);
  const double scale = (op == OP_HELMHOLTZ) ? dt * diffusion_coeff : dt;
  if (order == 1) {
    mg.helmholtz_lambda = 1 / scale;
    diffusion_set_rhs(mg, -1 / scale, 0.0);
  } else if (order == 2) {
    mg.helmholtz_lambda = 0.0;
    apply_op(mg, IRHS);
    mg.helmholtz_lambda = 2 / scale;
    diffusion_set_rhs(mg, -2 / scale, -1.0);
  } else fail("diffusion_solve: order should be 1 or 2");
  double res = 0.0;
  fas_fmg(mg, true, true, &res);
  int n;
  for (n = 1; n <= max_its; ++n) {
    if (res <= max_res) break;
    fas_vcycle(mg, false, 0, true, &res, true);
  }
  if (res_out) *res_out = res;
  if (n == max_its + 1) fail("diffusion_solve: no convergence");
  return n - 1;
}

#include "free_space.inc"

}  // namespace

// ===========================================================================
// Flat C API (ctypes).  All functions return 0 on success; on failure the
// message is available through orc_last_error (the reference would error stop).
// ===========================================================================
#define ORC_TRY(mgp, body) try { body; return 0; } catch (const OracleError& e) { (mgp)->err = e.msg; return 1; }

extern "C" {

void* orc_create() {
  Mg* mg = new Mg();
  std::memset(mg->box_size_lvl, 0, sizeof(mg->box_size_lvl));
  std::memset(mg->domain_size_lvl, 0, sizeof(mg->domain_size_lvl));
  std::memset(mg->dr, 0, sizeof(mg->dr));
  const char* names[N_TIMERS] = {"mg total V-cycle", "mg total FMG cycle", "mg smoother", "mg smoother g.c.",
                                 "mg coarse", "mg correct", "mg update coarse"};
  for (int t = 0; t < N_TIMERS; ++t) mg->timers[t].name = names[t];
  return mg;
}
void orc_destroy(void* h) { delete (Mg*)h; }
const char* orc_last_error(void* h) { return ((Mg*)h)->err.c_str(); }

int orc_set_int(void* h, const char* key, int v) {
  Mg* mg = (Mg*)h; std::string k(key);
  if (k == "operator_type") mg->operator_type = v;
  else if (k == "smoother_type") mg->smoother_type = v;
  else if (k == "n_extra_vars") mg->n_extra_vars = v;
  else if (k == "n_cycle_down") mg->n_cycle_down = v;
  else if (k == "n_cycle_up") mg->n_cycle_up = v;
  else if (k == "max_coarse_cycles") mg->max_coarse_cycles = v;
  else if (k == "n_cpu") mg->n_cpu = v;
  else if (k == "highest_lvl") mg->highest_lvl = v;
  else if (k == "coarsest_grid") { mg->coarsest_grid[0] = mg->coarsest_grid[1] = mg->coarsest_grid[2] = v; }
  else if (k == "subtract_mean") mg->subtract_mean = (v != 0);
  else { mg->err = "unknown int key " + k; return 1; }
  return 0;
}
int orc_get_int(void* h, const char* key) {
  Mg* mg = (Mg*)h; std::string k(key);
  if (k == "n_boxes") return mg->n_boxes;
  if (k == "lowest_lvl") return mg->lowest_lvl;
  if (k == "highest_lvl") return mg->highest_lvl;
  if (k == "first_normal_lvl") return mg->first_normal_lvl;
  if (k == "box_size") return mg->box_size;
  if (k == "n_extra_vars") return mg->n_extra_vars;
  if (k == "n_cpu") return mg->n_cpu;
  if (k == "subtract_mean") return mg->subtract_mean ? 1 : 0;
  if (k == "n_smoother_substeps") return mg->n_smoother_substeps;
  if (k == "last_coarse_iters") return mg->last_coarse_iters;
  if (k == "phi_bc_data_stored") return mg->phi_bc_data_stored ? 1 : 0;
  return -999999;
}
int orc_set_real(void* h, const char* key, double v) {
  Mg* mg = (Mg*)h; std::string k(key);
  if (k == "residual_coarse_abs") mg->residual_coarse_abs = v;
  else if (k == "residual_coarse_rel") mg->residual_coarse_rel = v;
  else if (k == "lambda") { if (v < 0) { mg->err = "set_lambda: lambda < 0 not allowed"; return 1; } mg->helmholtz_lambda = v; }
  else { mg->err = "unknown real key " + k; return 1; }
  return 0;
}
int orc_box_size_lvl(void* h, int lvl) { return ((Mg*)h)->bsl(lvl); }
void orc_dr_lvl(void* h, int lvl, double* out) { for (int d = 0; d < 3; ++d) out[d] = ((Mg*)h)->drl(lvl)[d]; }

int orc_set_bc(void* h, int nb1, int iv1, int bc_type, double bc_value) {
  Mg* mg = (Mg*)h; Bc& c = mg->bc[nb1 - 1][iv1 - 1];
  c.bc_type = bc_type; c.bc_value = bc_value; c.boundary_cond = nullptr; return 0;
}
int orc_set_bc_callback(void* h, int nb1, int iv1, bc_callback_t cb, void* user) {
  Mg* mg = (Mg*)h; Bc& c = mg->bc[nb1 - 1][iv1 - 1];
  c.boundary_cond = cb; c.user = user; return 0;
}

int orc_set_methods(void* h) { Mg* mg = (Mg*)h; ORC_TRY(mg, set_methods(*mg)); }
int orc_build_rectangle(void* h, const int* domain_size, int box_size, const double* dx, const double* r_min,
                        const int* periodic, int n_finer) {
  Mg* mg = (Mg*)h; bool per[3] = {periodic[0] != 0, periodic[1] != 0, periodic[2] != 0};
  ORC_TRY(mg, build_rectangle(*mg, domain_size, box_size, dx, r_min, per, n_finer));
}
int orc_add_children(void* h, int id) { Mg* mg = (Mg*)h; ORC_TRY(mg, add_children(*mg, id)); }
int orc_set_leaves_parents(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, set_leaves_parents(*mg, lvl)); }
int orc_set_next_level_ids(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, set_next_level_ids(*mg, lvl)); }
int orc_set_neighbors_lvl(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, set_neighbors_lvl(*mg, lvl)); }
int orc_set_refinement_boundaries(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, set_refinement_boundaries(*mg, lvl)); }
int orc_load_balance(void* h, int kind) {
  Mg* mg = (Mg*)h;
  ORC_TRY(mg, { if (kind == 0) load_balance_simple(*mg); else if (kind == 1) load_balance(*mg); else load_balance_parents(*mg); });
}
int orc_allocate_storage(void* h) { Mg* mg = (Mg*)h; ORC_TRY(mg, allocate_storage(*mg)); }

// level id lists: which = 0 ids, 1 leaves, 2 parents, 3 ref_bnds; returns count, copies into out if non-null
int orc_lvl_list(void* h, int lvl, int which, int* out) {
  Mg* mg = (Mg*)h; Lvl& L = mg->L(lvl);
  const std::vector<int>& v = which == 0 ? L.ids : which == 1 ? L.leaves : which == 2 ? L.parents : L.ref_bnds;
  if (out) std::copy(v.begin(), v.end(), out);
  return (int)v.size();
}
// box metadata: out[0]=rank, lvl, ix(3), parent, children(8), neighbors(6)  (20 ints)
void orc_box_info(void* h, int id, int* out, double* r_min) {
  Mg* mg = (Mg*)h; const Box& b = mg->boxes[id];
  int m = 0; out[m++] = b.rank; out[m++] = b.lvl;
  for (int d = 0; d < 3; ++d) out[m++] = b.ix[d];
  out[m++] = b.parent;
  for (int c = 0; c < 8; ++c) out[m++] = b.children[c];
  for (int n = 0; n < 6; ++n) out[m++] = b.neighbors[n];
  if (r_min) for (int d = 0; d < 3; ++d) r_min[d] = b.r_min[d];
}
// bulk tree export: per box 20 ints, for ids 1..n_boxes
void orc_tree_export(void* h, int* out) {
  Mg* mg = (Mg*)h;
  for (int id = 1; id <= mg->n_boxes; ++id) orc_box_info(h, id, out + (size_t)(id - 1) * 20, nullptr);
}
void orc_set_box_rank(void* h, int id, int rank) { ((Mg*)h)->boxes[id].rank = rank; }
int orc_update_lvl_info(void* h) { Mg* mg = (Mg*)h; ORC_TRY(mg, update_lvl_info(*mg)); }
double* orc_box_cc(void* h, int id) { return ((Mg*)h)->boxes[id].cc.data(); }

int orc_fill_ghost_cells_lvl(void* h, int lvl, int iv1) { Mg* mg = (Mg*)h; ORC_TRY(mg, fill_ghost_cells_lvl(*mg, lvl, iv1 - 1)); }
int orc_fill_ghost_cells(void* h, int iv1) {
  Mg* mg = (Mg*)h;
  ORC_TRY(mg, for (int lvl = mg->lowest_lvl; lvl <= mg->highest_lvl; ++lvl) fill_ghost_cells_lvl(*mg, lvl, iv1 - 1));
}
int orc_restrict_lvl(void* h, int iv1, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, restrict_lvl(*mg, iv1 - 1, lvl)); }
int orc_restrict(void* h, int iv1) {
  Mg* mg = (Mg*)h;
  ORC_TRY(mg, for (int lvl = mg->highest_lvl; lvl >= mg->lowest_lvl + 1; --lvl) restrict_lvl(*mg, iv1 - 1, lvl));
}
int orc_prolong(void* h, int lvl, int iv1, int iv_to1, int add) { Mg* mg = (Mg*)h; ORC_TRY(mg, prolong(*mg, lvl, iv1 - 1, iv_to1 - 1, add != 0)); }
int orc_apply_op(void* h, int i_out1) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); apply_op(*mg, i_out1 - 1); }); }
int orc_box_op(void* h, int id, int i_out1) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); box_op(*mg, id, mg->boxes[id].nc, i_out1 - 1); }); }
int orc_phi_bc_store(void* h) { Mg* mg = (Mg*)h; ORC_TRY(mg, phi_bc_store(*mg)); }
int orc_smooth_boxes(void* h, int lvl, int n_cycle) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); smooth_boxes(*mg, lvl, n_cycle); }); }
int orc_update_coarse(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); update_coarse(*mg, lvl); }); }
int orc_correct_children(void* h, int lvl) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); correct_children(*mg, lvl); }); }
int orc_max_residual_lvl(void* h, int lvl, double* res) { Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); *res = max_residual_lvl(*mg, lvl); }); }

// highest_lvl: pass has_highest = 0 for the reference's "not present"
int orc_fas_vcycle(void* h, int has_highest, int highest_lvl, int want_res, double* max_res, int standalone) {
  Mg* mg = (Mg*)h; ORC_TRY(mg, fas_vcycle(*mg, has_highest != 0, highest_lvl, want_res != 0, max_res, standalone != 0));
}
int orc_fas_fmg(void* h, int have_guess, int want_res, double* max_res) {
  Mg* mg = (Mg*)h; ORC_TRY(mg, fas_fmg(*mg, have_guess != 0, want_res != 0, max_res));
}
int orc_poisson_free_3d(void* h, int new_rhs, double max_fft_frac, int fmgcycle, int want_res, double* max_res) {
  Mg* mg = (Mg*)h; ORC_TRY(mg, { if (!mg->methods_set) set_methods(*mg); poisson_free_3d(*mg, new_rhs != 0, max_fft_frac, fmgcycle != 0, want_res != 0, max_res); });
}
int orc_free_space_fft_lvl(void* h) { Mg* mg = (Mg*)h; return mg->free_bc ? mg->free_bc->fft_lvl : -999; }
int orc_diffusion_set_rhs(void* h, double f1, double f2) { Mg* mg = (Mg*)h; ORC_TRY(mg, diffusion_set_rhs(*mg, f1, f2)); }
int orc_diffusion_solve(void* h, int op, double dt, double diffusion_coeff, int order, double max_res, double* res, int* n_vcycles) {
  Mg* mg = (Mg*)h; ORC_TRY(mg, { int n = diffusion_solve(*mg, op, dt, diffusion_coeff, order, max_res, res); if (n_vcycles) *n_vcycles = n; });
}
double orc_timer(void* h, int t) { return ((Mg*)h)->timers[t].t; }
int orc_omp_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_omp_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

}  // extern "C"
