// m_octree_mg.cpp -- C++ host mirror of the reference's Fortran API (see m_octree_mg.hpp) plus a
// flat C wrapper (omgh_*) used by the Python harness.  Tree construction, load balancing and the
// host data model follow the reference (file:line cited per routine); the cycle itself is the GPU
// library behind include/omg_gpu.h.
#include "m_octree_mg.hpp"

#include <algorithm>
#include <cstring>
#include <map>

namespace omg {

namespace {
// topology tables, src/m_data_structures.f90:150-191 (values 1-based as in the reference)
const int mg_child_dix[8][3] = {{0,0,0},{1,0,0},{0,1,0},{1,1,0},{0,0,1},{1,0,1},{0,1,1},{1,1,1}};
const int mg_child_rev[3][8] = {{2,1,4,3,6,5,8,7},{3,4,1,2,7,8,5,6},{5,6,7,8,1,2,3,4}};
const int mg_child_adj_nb[6][4] = {{1,3,5,7},{2,4,6,8},{1,2,5,6},{3,4,7,8},{1,2,3,4},{5,6,7,8}};
const bool mg_neighb_low[6] = {true,false,true,false,true,false};
const int mg_neighb_rev[6] = {2,1,4,3,6,5};
const int mg_neighb_dim[6] = {0,0,1,1,2,2};

[[noreturn]] void stop(const std::string& msg) { throw mg_error(msg); }
void gpu_check(mg_t& mg, int rc, const char* where) {
  if (rc != 0) stop(std::string(where) + ": " + omg_last_error(mg.gpu));
}
void need_gpu(const mg_t& mg) { if (!mg.gpu) stop("device storage not allocated: call mg_allocate_storage first"); }
}  // namespace

mg_t::mg_t() {
  std::memset(box_size_lvl, 0, sizeof(box_size_lvl));
  std::memset(domain_size_lvl, 0, sizeof(domain_size_lvl));
  std::memset(dr, 0, sizeof(dr));
}
mg_t::~mg_t() { if (gpu) omg_destroy(gpu); }

bool mg_has_children(const mg_box_t& box) { return box.children[0] != mg_no_box; }

int mg_ix_to_ichild(const int ix[3]) { return 8 - 4 * (ix[2] & 1) - 2 * (ix[1] & 1) - (ix[0] & 1); }

void mg_get_child_offset(const mg_t& mg, int id, int off[3]) {
  const mg_box_t& b = mg.boxes[id];
  for (int d = 0; d < 3; ++d)
    off[d] = (b.lvl <= mg.first_normal_lvl) ? 0 : ((b.ix[d] - 1) & 1) * (mg.box_size >> 1);
}

int64_t mg_number_of_unknowns(const mg_t& mg) {
  int64_t n = 0;
  for (int lvl = mg.first_normal_lvl; lvl <= mg.highest_lvl; ++lvl) n += (int64_t)mg.L(lvl).leaves.size();
  return n * (int64_t)mg.box_size * mg.box_size * mg.box_size;
}

// src/m_data_structures.f90:495-539
void mg_get_face_coords(const mg_box_t& box, int nb, int nc, double* x) {
  const int d = mg_neighb_dim[nb - 1];
  const int ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
  double rmin[3] = {box.r_min[0], box.r_min[1], box.r_min[2]};
  if (!mg_neighb_low[nb - 1]) rmin[d] = rmin[d] + box.dr[d] * nc;
  const size_t n2 = (size_t)nc * nc;
  for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
    const size_t q = (size_t)(j - 1) * nc + (i - 1);
    x[0 * n2 + q] = rmin[0]; x[1 * n2 + q] = rmin[1]; x[2 * n2 + q] = rmin[2];
    x[ta * n2 + q] = x[ta * n2 + q] + (i - 0.5) * box.dr[ta];
    x[tb * n2 + q] = x[tb * n2 + q] + (j - 0.5) * box.dr[tb];
  }
}

void mg_comm_init(mg_t& mg, int my_rank, int n_cpu) { mg.my_rank = my_rank; mg.n_cpu = n_cpu; }

// ------------------------------------------------------------------------------------------------
// m_build_tree
// ------------------------------------------------------------------------------------------------
void mg_set_leaves_parents(std::vector<mg_box_t>& boxes, mg_lvl_t& level) {      // :258-293
  level.parents.clear(); level.leaves.clear();
  for (int id : level.ids) (mg_has_children(boxes[id]) ? level.parents : level.leaves).push_back(id);
}

void mg_set_next_level_ids(mg_t& mg, int lvl) {                                   // :187-216
  std::vector<int>& next = mg.L(lvl + 1).ids;
  next.clear();
  const bool octree = mg.box_size_lvl[lvl + 1 - mg_lvl_lo] == mg.box_size_lvl[lvl - mg_lvl_lo];
  for (int id : mg.L(lvl).parents) {
    if (octree) next.insert(next.end(), mg.boxes[id].children, mg.boxes[id].children + 8);
    else next.push_back(mg.boxes[id].children[0]);
  }
}

static int find_neighb(const std::vector<mg_box_t>& boxes, int id, int nb) {       // :236-255
  int p_id = boxes[id].parent;
  const int c_ix = mg_ix_to_ichild(boxes[id].ix), d = mg_neighb_dim[nb - 1];
  const bool child_low = mg_child_dix[c_ix - 1][d] == 0;
  if (child_low == mg_neighb_low[nb - 1]) p_id = boxes[p_id].neighbors[nb - 1];
  if (p_id <= mg_no_box) return p_id;
  return boxes[p_id].children[mg_child_rev[d][c_ix - 1] - 1];
}

void mg_set_neighbors_lvl(mg_t& mg, int lvl) {                                    // :176-185, :218-233
  for (int id : mg.L(lvl).ids)
    for (int nb = 1; nb <= 6; ++nb)
      if (mg.boxes[id].neighbors[nb - 1] == mg_no_box) {
        const int nb_id = find_neighb(mg.boxes, id, nb);
        if (nb_id > mg_no_box) {
          mg.boxes[id].neighbors[nb - 1] = nb_id;
          mg.boxes[nb_id].neighbors[mg_neighb_rev[nb - 1] - 1] = id;
        }
      }
}

void mg_set_refinement_boundaries(std::vector<mg_box_t>& boxes, mg_lvl_t& level) { // :296-328
  level.ref_bnds.clear();
  if (level.parents.empty()) return;
  for (int id : level.leaves)
    for (int nb = 0; nb < 6; ++nb) {
      const int nb_id = boxes[id].neighbors[nb];
      if (nb_id > mg_no_box && mg_has_children(boxes[nb_id])) { level.ref_bnds.push_back(id); break; }
    }
}

void mg_add_children(mg_t& mg, int id) {                                          // :330-367
  if (mg.n_boxes + mg_num_children > (int)mg.boxes.size() - 1) stop("mg_add_children: not enough space");
  const int first = mg.n_boxes + 1;
  mg.n_boxes += mg_num_children;
  const int lvl = mg.boxes[id].lvl + 1;
  for (int i = 0; i < 8; ++i) {
    mg_box_t& c = mg.boxes[first + i];
    const mg_box_t& p = mg.boxes[id];
    c = mg_box_t();
    c.rank = p.rank; c.id = first + i; c.lvl = lvl; c.parent = id;
    for (int d = 0; d < 3; ++d) {
      c.ix[d] = 2 * p.ix[d] - 1 + mg_child_dix[i][d];
      c.r_min[d] = p.r_min[d] + mg.dr[lvl - mg_lvl_lo][d] * mg_child_dix[i][d] * mg.box_size;
      c.dr[d] = mg.dr[lvl - mg_lvl_lo][d];
    }
    mg.boxes[id].children[i] = first + i;
  }
  for (int nb = 0; nb < 6; ++nb) {
    const int pn = mg.boxes[id].neighbors[nb];
    if (pn < mg_no_box)
      for (int n = 0; n < 4; ++n) mg.boxes[first + mg_child_adj_nb[nb][n] - 1].neighbors[nb] = pn;
  }
}

static void add_single_child(mg_t& mg, int id, int n_boxes_lvl) {                  // :369-393
  const int c_id = ++mg.n_boxes;
  mg_box_t& p = mg.boxes[id];
  mg_box_t& c = mg.boxes[c_id];
  c = mg_box_t();
  p.children[0] = c_id;
  c.rank = p.rank; c.id = c_id; c.lvl = p.lvl + 1; c.parent = id;
  for (int d = 0; d < 3; ++d) { c.ix[d] = p.ix[d]; c.r_min[d] = p.r_min[d]; c.dr[d] = mg.dr[c.lvl - mg_lvl_lo][d]; }
  for (int n = 0; n < 6; ++n)
    c.neighbors[n] = (p.neighbors[n] == mg_physical_boundary) ? p.neighbors[n] : p.neighbors[n] + n_boxes_lvl;
}

void mg_build_rectangle(mg_t& mg, const int domain_size[3], int box_size, const double dx[3], const double r_min[3],
                        const bool periodic[3], int n_finer) {                     // :18-174
  if (box_size % 2 != 0) stop("box_size should be even");
  for (int d = 0; d < 3; ++d) if (domain_size[d] % box_size != 0) stop("box_size does not divide domain_size");
  if (periodic[0] && periodic[1] && periodic[2]) mg.subtract_mean = true;

  auto bsl = [&](int lvl) -> int& { return mg.box_size_lvl[lvl - mg_lvl_lo]; };
  int per_dim[OMG_NLVL][3];
  std::memset(per_dim, 0, sizeof(per_dim));
  int nx[3];
  mg.box_size = box_size; bsl(1) = box_size; mg.first_normal_lvl = 1;
  for (int d = 0; d < 3; ++d) {
    nx[d] = domain_size[d];
    mg.domain_size_lvl[1 - mg_lvl_lo][d] = nx[d];
    mg.dr[1 - mg_lvl_lo][d] = dx[d];
    mg.r_min[d] = r_min[d]; mg.periodic[d] = periodic[d];
    per_dim[1 - mg_lvl_lo][d] = nx[d] / box_size;
  }
  int lvl = 1;
  for (; lvl > mg_lvl_lo; --lvl) {
    bool done = false, halve_boxes = true;
    for (int d = 0; d < 3; ++d) {
      done = done || nx[d] % 2 == 1 || nx[d] == mg.coarsest_grid[d] ||
             (bsl(lvl) == mg.coarsest_grid[d] && mg.smoother_type == mg_smoother_gs);
      halve_boxes = halve_boxes && (nx[d] / bsl(lvl)) % 2 == 0;
    }
    if (done) break;
    bsl(lvl - 1) = halve_boxes ? bsl(lvl) : bsl(lvl) / 2;
    if (halve_boxes) mg.first_normal_lvl = lvl - 1;
    for (int d = 0; d < 3; ++d) {
      per_dim[lvl - 1 - mg_lvl_lo][d] = halve_boxes ? per_dim[lvl - mg_lvl_lo][d] / 2 : per_dim[lvl - mg_lvl_lo][d];
      mg.dr[lvl - 1 - mg_lvl_lo][d] = mg.dr[lvl - mg_lvl_lo][d] * 2;
      nx[d] /= 2;
      mg.domain_size_lvl[lvl - 1 - mg_lvl_lo][d] = nx[d];
    }
  }
  mg.lowest_lvl = lvl;
  mg.highest_lvl = 1;
  for (int l = 2; l <= mg_lvl_hi; ++l) {
    bsl(l) = box_size;
    for (int d = 0; d < 3; ++d) {
      mg.dr[l - mg_lvl_lo][d] = mg.dr[l - 1 - mg_lvl_lo][d] * 0.5;
      mg.domain_size_lvl[l - mg_lvl_lo][d] = 2 * mg.domain_size_lvl[l - 1 - mg_lvl_lo][d];
    }
  }
  size_t total = (size_t)n_finer;
  for (int l = mg_lvl_lo; l <= 1; ++l) total += (size_t)per_dim[l - mg_lvl_lo][0] * per_dim[l - mg_lvl_lo][1] * per_dim[l - mg_lvl_lo][2];
  mg.boxes.assign(total + 1, mg_box_t());
  mg.n_boxes = 0;
  for (auto& L : mg.lvls) L = mg_lvl_t();

  const int* nb3 = per_dim[mg.lowest_lvl - mg_lvl_lo];
  const int stride[3] = {1, nb3[0], nb3[0] * nb3[1]};
  const int lo = mg.lowest_lvl;
  for (int k = 1; k <= nb3[2]; ++k) for (int j = 1; j <= nb3[1]; ++j) for (int i = 1; i <= nb3[0]; ++i) {
    const int n = ++mg.n_boxes;
    mg_box_t& b = mg.boxes[n];
    b.rank = 0; b.id = n; b.lvl = lo; b.parent = mg_no_box;
    const int ijk[3] = {i, j, k};
    for (int d = 0; d < 3; ++d) {
      b.ix[d] = ijk[d];
      b.r_min[d] = r_min[d] + (ijk[d] - 1) * bsl(lo) * mg.dr[lo - mg_lvl_lo][d];
      b.dr[d] = mg.dr[lo - mg_lvl_lo][d];
      b.neighbors[2 * d] = n - stride[d];
      b.neighbors[2 * d + 1] = n + stride[d];
      const int wrap = (nb3[d] - 1) * stride[d];
      if (ijk[d] == 1) b.neighbors[2 * d] = periodic[d] ? n + wrap : mg_physical_boundary;
      if (ijk[d] == nb3[d]) b.neighbors[2 * d + 1] = periodic[d] ? n - wrap : mg_physical_boundary;
    }
    mg.L(lo).ids.push_back(n);
  }
  for (int l = lo; l <= 0; ++l) {
    const std::vector<int> ids = mg.L(l).ids;
    if (bsl(l + 1) == bsl(l)) {
      for (int id : ids) mg_add_children(mg, id);
      mg_set_leaves_parents(mg.boxes, mg.L(l));
      mg_set_next_level_ids(mg, l);
      mg_set_neighbors_lvl(mg, l + 1);
    } else {
      for (int id : ids) add_single_child(mg, id, (int)ids.size());
      mg_set_leaves_parents(mg.boxes, mg.L(l));
      mg_set_next_level_ids(mg, l);
    }
  }
  mg_set_leaves_parents(mg.boxes, mg.L(1));
  for (int l = lo; l <= 1; ++l) mg.L(l).ref_bnds.clear();
  mg.tree_created = true;
}

// ------------------------------------------------------------------------------------------------
// m_load_balance
// ------------------------------------------------------------------------------------------------
static int most_popular(const int* list, int n, const std::vector<int>& work) {   // :197-221
  int best_count = 0, best_work = 0, res = -1;
  for (int i = 0; i < n; ++i) {
    int cnt = 0;
    for (int j = 0; j < n; ++j) cnt += (list[j] == list[i]);
    const int w = work[list[i]];
    if (cnt > best_count || (cnt == best_count && w < best_work)) { best_count = cnt; best_work = w; res = list[i]; }
  }
  return res;
}

static void update_lvl_info(mg_t& mg) {                                           // :223-235
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    mg_lvl_t& L = mg.L(lvl);
    auto mine = [&](const std::vector<int>& all, std::vector<int>& out) {
      out.clear();
      for (int id : all) if (mg.boxes[id].rank == mg.my_rank) out.push_back(id);
    };
    mine(L.ids, L.my_ids); mine(L.leaves, L.my_leaves); mine(L.parents, L.my_parents); mine(L.ref_bnds, L.my_ref_bnds);
  }
}

static void assign_coarse_rank(mg_t& mg, int single_cpu_lvl, const std::vector<int>& my_work) {
  int coarse_rank = 0;
  if (single_cpu_lvl < mg.highest_lvl) {
    std::vector<int> ranks;
    for (int id : mg.L(single_cpu_lvl + 1).ids) ranks.push_back(mg.boxes[id].rank);
    coarse_rank = most_popular(ranks.data(), (int)ranks.size(), my_work);
  }
  for (int lvl = mg.lowest_lvl; lvl <= single_cpu_lvl; ++lvl)
    for (int id : mg.L(lvl).ids) mg.boxes[id].rank = coarse_rank;
  // B200 extension (off by default = the reference's partition): coarse levels that no longer fill the
  // GPUs collapse onto the coarse rank -- the levels above single_cpu_lvl, bottom up, while they hold fewer
  // than coarse_collapse_boxes boxes per rank.  Results do not depend on the partition.
  if (mg.coarse_collapse_boxes > 0 && mg.n_cpu > 1)
    for (int lvl = single_cpu_lvl + 1; lvl < mg.highest_lvl; ++lvl) {
      if ((long)mg.L(lvl).ids.size() >= (long)mg.coarse_collapse_boxes * mg.n_cpu) break;
      for (int id : mg.L(lvl).ids) mg.boxes[id].rank = coarse_rank;
    }
}

void mg_load_balance_simple(mg_t& mg) {                                           // :22-63
  const int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  for (int lvl = mg.lowest_lvl; lvl <= single_cpu_lvl; ++lvl) for (int id : mg.L(lvl).ids) mg.boxes[id].rank = 0;
  for (int lvl = single_cpu_lvl + 1; lvl <= mg.highest_lvl; ++lvl) {
    int work_left = (int)mg.L(lvl).ids.size(), my_work = 0, i_cpu = 0;
    for (int id : mg.L(lvl).ids) {
      if ((mg.n_cpu - i_cpu - 1) * my_work >= work_left) { ++i_cpu; my_work = 0; }
      ++my_work; --work_left;
      mg.boxes[id].rank = i_cpu;
    }
  }
  update_lvl_info(mg);
}

void mg_load_balance(mg_t& mg) {                                                  // :71-136
  const int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  std::vector<int> my_work(mg.n_cpu + 1, 0);
  for (int lvl = mg.highest_lvl; lvl > single_cpu_lvl; --lvl) {
    std::fill(my_work.begin(), my_work.end(), 0);
    for (int id : mg.L(lvl).parents) {
      int c_ranks[8];
      for (int c = 0; c < 8; ++c) c_ranks[c] = mg.boxes[mg.boxes[id].children[c]].rank;
      const int i_cpu = most_popular(c_ranks, 8, my_work);
      mg.boxes[id].rank = i_cpu;
      ++my_work[i_cpu];
    }
    int work_left = (int)mg.L(lvl).leaves.size(), i_cpu = 0;
    for (int id : mg.L(lvl).leaves) {
      int rest = 0;
      for (int r = i_cpu + 1; r <= mg.n_cpu; ++r) rest += my_work[r];
      if ((mg.n_cpu - i_cpu - 1) * my_work[i_cpu] >= work_left + rest) ++i_cpu;
      ++my_work[i_cpu]; --work_left;
      mg.boxes[id].rank = i_cpu;
    }
  }
  assign_coarse_rank(mg, single_cpu_lvl, my_work);
  update_lvl_info(mg);
}

void mg_load_balance_parents(mg_t& mg) {                                          // :140-193
  const int single_cpu_lvl = std::max(mg.first_normal_lvl - 1, mg.lowest_lvl);
  std::vector<int> my_work(mg.n_cpu + 1, 0);
  for (int lvl = mg.highest_lvl - 1; lvl > single_cpu_lvl; --lvl) {
    std::fill(my_work.begin(), my_work.end(), 0);
    for (int id : mg.L(lvl).leaves) ++my_work[mg.boxes[id].rank];
    for (int id : mg.L(lvl).parents) {
      int c_ranks[8];
      for (int c = 0; c < 8; ++c) c_ranks[c] = mg.boxes[mg.boxes[id].children[c]].rank;
      const int i_cpu = most_popular(c_ranks, 8, my_work);
      mg.boxes[id].rank = i_cpu;
      ++my_work[i_cpu];
    }
  }
  assign_coarse_rank(mg, single_cpu_lvl, my_work);
  update_lvl_info(mg);
}

// ------------------------------------------------------------------------------------------------
// m_allocate_storage: host cc arrays for owned boxes + the device mirror behind the C ABI
// ------------------------------------------------------------------------------------------------
static void push_params(mg_t& mg);
// flatten type(mg_t) for the C ABI and hand it to omg_set_tree (update = false) or omg_update_tree (update = true)
static void hand_over_tree(mg_t& mg, bool update, const char* where) {
  const int n = mg.n_boxes;
  std::vector<int32_t> rank(n), lvl(n), ix(3 * (size_t)n), parent(n), children(8 * (size_t)n), neighbors(6 * (size_t)n);
  for (int id = 1; id <= n; ++id) {
    const mg_box_t& b = mg.boxes[id];
    rank[id - 1] = b.rank; lvl[id - 1] = b.lvl; parent[id - 1] = b.parent;
    for (int d = 0; d < 3; ++d) ix[(size_t)(id - 1) * 3 + d] = b.ix[d];
    for (int c = 0; c < 8; ++c) children[(size_t)(id - 1) * 8 + c] = b.children[c];
    for (int f = 0; f < 6; ++f) neighbors[(size_t)(id - 1) * 6 + f] = b.neighbors[f];
  }
  std::vector<int32_t> lvl_ptr(OMG_NLVL + 1, 0), lvl_ids;
  for (int l = mg_lvl_lo; l <= mg_lvl_hi; ++l) {
    if (l >= mg.lowest_lvl && l <= mg.highest_lvl) lvl_ids.insert(lvl_ids.end(), mg.L(l).ids.begin(), mg.L(l).ids.end());
    lvl_ptr[l - mg_lvl_lo + 1] = (int32_t)lvl_ids.size();
  }
  omg_tree_desc t{};
  t.n_boxes = n; t.box_size = mg.box_size; t.lowest_lvl = mg.lowest_lvl; t.highest_lvl = mg.highest_lvl;
  t.first_normal_lvl = mg.first_normal_lvl; t.my_rank = mg.my_rank; t.n_cpu = mg.n_cpu; t.n_extra_vars = mg.n_extra_vars;
  for (int d = 0; d < 3; ++d) t.periodic[d] = mg.periodic[d];
  t.box_size_lvl = mg.box_size_lvl; t.dr = &mg.dr[0][0];
  t.box_rank = rank.data(); t.box_lvl = lvl.data(); t.box_ix = ix.data(); t.box_parent = parent.data();
  t.box_children = children.data(); t.box_neighbors = neighbors.data();
  t.lvl_ptr = lvl_ptr.data(); t.lvl_ids = lvl_ids.data();
  if (!mg.gpu) {
    const int rc = omg_create(mg.device, &mg.gpu);
    if (rc != 0 || !mg.gpu) { mg.gpu = nullptr; stop("mg_allocate_storage: no CUDA device available (there is no CPU fallback)"); }
  }
  gpu_check(mg, update ? omg_update_tree(mg.gpu, &t) : omg_set_tree(mg.gpu, &t), where);
}

void mg_allocate_storage(mg_t& mg) {                                              // :51-99
  if (!mg.tree_created) stop("allocate_storage: tree is not yet created");
  if (mg.is_allocated) stop("allocate_storage: tree is already allocated");
  if (mg.my_rank < 0 || mg.n_cpu < 1) mg_comm_init(mg);
  update_lvl_info(mg);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo];
    if (mg.host_cc) for (int id : mg.L(lvl).my_ids) mg.boxes[id].cc.assign((size_t)(nc + 2) * (nc + 2) * (nc + 2) * mg.nvar(), 0.0);
  }
  hand_over_tree(mg, false, "mg_allocate_storage");
  mg.bc_dirty = true; mg.eps_dirty = true;
  mg.is_allocated = true;
}

// On-device regrid: the host has changed the tree of an allocated mg_t (mg_add_children + mg_set_neighbors_lvl +
// mg_set_refinement_boundaries, src/m_build_tree.f90:296-367; mg_load_balance*, src/m_load_balance.f90:22-235).  Boxes that
// exist before and after keep their data on the device (omg_update_tree: local copy, or old owner -> new owner); new boxes
// start from zero like freshly allocated storage.  The reference's own route is mg_deallocate_storage + a rebuild + copying
// every variable in again (coupling_amrvac).  Host cc arrays follow: allocated for boxes this rank owns now, released for
// the others, and in drop-in mode refreshed from the device (boxes that arrived from another rank exist only there).
void mg_gpu_update_tree(mg_t& mg) {
  if (!mg.is_allocated) stop("mg_gpu_update_tree: tree is not allocated");
  need_gpu(mg);
  const bool drop_in = !mg.gpu_resident && mg.host_cc;
  if (drop_in) {
    // drop-in mode: the host arrays are the truth.  Everything this rank held so far (= every box with a cc array) goes to
    // the device first, so that boxes which change owner take ALL their variables along
    std::vector<int32_t> ids;
    for (int id = 1; id <= mg.n_boxes; ++id) if (!mg.boxes[id].cc.empty()) ids.push_back(id);
    std::vector<const double*> ptrs(ids.size());
    for (int iv = 1; iv <= mg.nvar(); ++iv) {
      for (size_t q = 0; q < ids.size(); ++q) {
        const int n2 = mg.box_size_lvl[mg.boxes[ids[q]].lvl - mg_lvl_lo] + 2;
        ptrs[q] = mg.boxes[ids[q]].cc.data() + (size_t)(iv - 1) * n2 * n2 * n2;
      }
      gpu_check(mg, omg_push_var(mg.gpu, iv, (int)ids.size(), ids.data(), ptrs.data()), "mg_gpu_update_tree");
    }
  }
  update_lvl_info(mg);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo];
    const size_t want = (size_t)(nc + 2) * (nc + 2) * (nc + 2) * mg.nvar();
    for (int id : mg.L(lvl).ids) {
      mg_box_t& box = mg.boxes[id];
      if (box.rank == mg.my_rank) { if (mg.host_cc && box.cc.size() != want) box.cc.assign(want, 0.0); }
      else if (!box.cc.empty()) std::vector<double>().swap(box.cc);
    }
  }
  hand_over_tree(mg, true, "mg_gpu_update_tree");
  mg.bc_dirty = true; mg.eps_dirty = true;
  mg.free_bc_initialized = false;
  if (drop_in) {
    push_params(mg);
    for (int iv = 1; iv <= mg.nvar(); ++iv) mg_gpu_pull(mg, iv, iv == mg_iphi || iv >= mg_iveps);
  }
}

void mg_deallocate_storage(mg_t& mg) {                                            // :14-47
  if (!mg.is_allocated) stop("deallocate_storage: tree is not allocated");
  if (mg.gpu) { omg_destroy(mg.gpu); mg.gpu = nullptr; }
  mg.boxes.clear();
  for (auto& L : mg.lvls) L = mg_lvl_t();
  mg.is_allocated = false; mg.n_boxes = 0; mg.phi_bc_data_stored = false;
}

// ------------------------------------------------------------------------------------------------
// m_multigrid
// ------------------------------------------------------------------------------------------------
void mg_set_methods(mg_t& mg) {                                                   // m_multigrid.f90:27-60
  const bool all_periodic = mg.periodic[0] && mg.periodic[1] && mg.periodic[2];
  auto neumann0 = [&](int iv) { for (int nb = 0; nb < 6; ++nb) { mg.bc[nb][iv - 1].bc_type = mg_bc_neumann; mg.bc[nb][iv - 1].bc_value = 0.0; } };
  auto extra = [&](int need, const char* who) {
    if (mg.n_extra_vars == 0 && mg.is_allocated) stop(std::string(who) + "_set_methods: mg%n_extra_vars == 0");
    mg.n_extra_vars = std::max(need, mg.n_extra_vars);
  };
  if (mg.geometry_type != mg_cartesian) stop("set_methods: unsupported geometry");
  switch (mg.operator_type) {
    case mg_laplacian: if (all_periodic) mg.subtract_mean = true; break;                              // m_laplacian.f90:13-49
    case mg_vlaplacian: extra(1, "vlaplacian"); mg.subtract_mean = false; neumann0(mg_iveps); break;   // m_vlaplacian.f90:14-48
    case mg_helmholtz: mg.subtract_mean = false; break;                                               // m_helmholtz.f90:18-37
    case mg_vhelmholtz: extra(1, "vhelmholtz"); mg.subtract_mean = false; neumann0(mg_iveps); break;   // m_vhelmholtz.f90:19-49
    case mg_ahelmholtz: extra(3, "ahelmholtz"); neumann0(mg_iveps1); neumann0(mg_iveps2); neumann0(mg_iveps3); break;  // m_ahelmholtz.f90:19-57
    default: stop("mg_set_methods: unknown operator");
  }
  if (mg.smoother_type != mg_smoother_gs && mg.smoother_type != mg_smoother_gsrb) stop("set_methods: unsupported smoother type");
  mg.n_smoother_substeps = (mg.smoother_type == mg_smoother_gsrb) ? 2 : 1;
  mg.methods_set = true;
}

void helmholtz_set_lambda(mg_t& mg, double lambda) {                              // m_helmholtz.f90:39-46
  if (lambda < 0) stop("helmholtz_set_lambda: lambda < 0 not allowed");
  mg.lambda = lambda;
}

static void push_params(mg_t& mg) {
  need_gpu(mg);
  if (!mg.methods_set) mg_set_methods(mg);                                        // check_methods :62-70
  omg_params p{};
  p.operator_type = mg.operator_type; p.smoother_type = mg.smoother_type; p.geometry_type = mg.geometry_type;
  p.n_cycle_down = mg.n_cycle_down; p.n_cycle_up = mg.n_cycle_up; p.max_coarse_cycles = mg.max_coarse_cycles;
  p.subtract_mean = mg.subtract_mean ? 1 : 0;
  p.residual_coarse_abs = mg.residual_coarse_abs; p.residual_coarse_rel = mg.residual_coarse_rel;
  p.lambda = mg.lambda;
  gpu_check(mg, omg_set_params(mg.gpu, &p), "mg_set_methods");
  for (int iv = 1; iv <= mg.nvar(); ++iv)
    for (int nb = 1; nb <= 6; ++nb)
      gpu_check(mg, omg_set_bc_const(mg.gpu, iv, nb, mg.bc[nb - 1][iv - 1].bc_type, mg.bc[nb - 1][iv - 1].bc_value), "bc");
  if (mg.bc_dirty) mg_gpu_push_bc(mg);
}

static void plane_from_ghosts(const mg_box_t& box, int nc, int iv, int nb, double* plane) {   // box_get_gc m_ghost_cells.f90:579-620
  const int n2 = nc + 2, d = mg_neighb_dim[nb - 1], g = mg_neighb_low[nb - 1] ? 0 : nc + 1;
  const double* v = box.cc.data() + (size_t)(iv - 1) * n2 * n2 * n2;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    const int i = d == 0 ? g : a, j = d == 1 ? g : (d == 0 ? a : b), k = d == 2 ? g : b;
    plane[(size_t)(b - 1) * nc + (a - 1)] = v[((size_t)k * n2 + j) * n2 + i];
  }
}
static void plane_to_ghosts(mg_box_t& box, int nc, int iv, int nb, const double* plane) {     // box_set_gc :622-663
  const int n2 = nc + 2, d = mg_neighb_dim[nb - 1], g = mg_neighb_low[nb - 1] ? 0 : nc + 1;
  double* v = box.cc.data() + (size_t)(iv - 1) * n2 * n2 * n2;
  for (int b = 1; b <= nc; ++b) for (int a = 1; a <= nc; ++a) {
    const int i = d == 0 ? g : a, j = d == 1 ? g : (d == 0 ? a : b), k = d == 2 ? g : b;
    v[((size_t)k * n2 + j) * n2 + i] = plane[(size_t)(b - 1) * nc + (a - 1)];
  }
}

// Boundary planes of phi: what set_ghost_cells (m_ghost_cells.f90:262-279) would obtain on every
// ghost fill -- stored planes (phi_bc_data_stored), callback values, or the constants.
void mg_gpu_push_bc(mg_t& mg) {
  need_gpu(mg);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo];
    std::vector<double> plane((size_t)nc * nc);
    for (int id : mg.L(lvl).my_ids) {
      mg_box_t& box = mg.boxes[id];
      for (int nb = 1; nb <= 6; ++nb) {
        const int nb_id = box.neighbors[nb - 1];
        if (nb_id >= mg_no_box) continue;
        int bc_type;
        const mg_bc_t& c = mg.bc[nb - 1][mg_iphi - 1];
        if (mg.phi_bc_data_stored) { plane_from_ghosts(box, nc, mg_irhs, nb, plane.data()); bc_type = nb_id; }
        else if (c.boundary_cond) c.boundary_cond(box, nc, mg_iphi, nb, &bc_type, plane.data(), c.user);
        else continue;   // constants: omg_set_bc_const
        gpu_check(mg, omg_set_bc_plane(mg.gpu, id, nb, bc_type, plane.data()), "mg_gpu_push_bc");
      }
    }
  }
  mg.bc_dirty = false;
}

void mg_phi_bc_store(mg_t& mg) {                                                  // m_ghost_cells.f90:66-117
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo];
    std::vector<double> plane((size_t)nc * nc);
    for (int id : mg.L(lvl).my_ids) {
      mg_box_t& box = mg.boxes[id];
      for (int nb = 1; nb <= 6; ++nb) {
        if (box.neighbors[nb - 1] >= mg_no_box) continue;
        int bc_type;
        const mg_bc_t& c = mg.bc[nb - 1][mg_iphi - 1];
        if (c.boundary_cond) c.boundary_cond(box, nc, mg_iphi, nb, &bc_type, plane.data(), c.user);
        else { bc_type = c.bc_type; std::fill(plane.begin(), plane.end(), c.bc_value); }
        box.neighbors[nb - 1] = bc_type;
        plane_to_ghosts(box, nc, mg_irhs, nb, plane.data());
      }
    }
  }
  mg.phi_bc_data_stored = true;
  mg.bc_dirty = true;
}

static void transfer(mg_t& mg, int iv, bool push, bool with_ghosts, int lvl_lo, int lvl_hi) {
  need_gpu(mg);
  std::vector<int32_t> ids;
  std::vector<const double*> cp;
  std::vector<double*> p;
  for (int lvl = std::max(lvl_lo, mg.lowest_lvl); lvl <= std::min(lvl_hi, mg.highest_lvl); ++lvl) {
    const int n2 = mg.box_size_lvl[lvl - mg_lvl_lo] + 2;
    for (int id : mg.L(lvl).my_ids) {
      double* v = mg.boxes[id].cc.data() + (size_t)(iv - 1) * n2 * n2 * n2;
      ids.push_back(id); cp.push_back(v); p.push_back(v);
    }
  }
  if (push) gpu_check(mg, omg_push_var(mg.gpu, iv, (int)ids.size(), ids.data(), cp.data()), "mg_gpu_push");
  else gpu_check(mg, omg_pull_var(mg.gpu, iv, (int)ids.size(), ids.data(), p.data(), with_ghosts ? 1 : 0), "mg_gpu_pull");
}
void mg_gpu_push(mg_t& mg, int iv) { transfer(mg, iv, true, false, mg_lvl_lo, mg_lvl_hi); }
void mg_gpu_pull(mg_t& mg, int iv, bool with_ghosts) { transfer(mg, iv, false, with_ghosts, mg_lvl_lo, mg_lvl_hi); }

void mg_gpu_push_params(mg_t& mg) { push_params(mg); }

static int n_eps_vars(const mg_t& mg) {
  return (mg.operator_type == mg_vlaplacian || mg.operator_type == mg_vhelmholtz) ? 1 : (mg.operator_type == mg_ahelmholtz ? 3 : 0);
}
static void sync_in(mg_t& mg) {
  push_params(mg);
  if (mg.gpu_resident) return;
  mg_gpu_push(mg, mg_iphi); mg_gpu_push(mg, mg_irhs);
  for (int e = 0; e < n_eps_vars(mg); ++e) mg_gpu_push(mg, mg_iveps + e);
}
static void sync_out(mg_t& mg) {
  if (mg.gpu_resident) return;
  mg_gpu_pull(mg, mg_iphi, true); mg_gpu_pull(mg, mg_irhs, false); mg_gpu_pull(mg, mg_iold, true); mg_gpu_pull(mg, mg_ires, false);
}

void mg_fas_fmg(mg_t& mg, bool have_guess, double* max_res) {                      // m_multigrid.f90:84-147
  sync_in(mg);
  gpu_check(mg, omg_fas_fmg(mg.gpu, have_guess ? 1 : 0, max_res ? 1 : 0, max_res), "mg_fas_fmg");
  sync_out(mg);
}

void mg_fas_vcycle(mg_t& mg, const int* highest_lvl, double* max_res, bool standalone) {   // :150-243
  sync_in(mg);
  gpu_check(mg, omg_fas_vcycle(mg.gpu, highest_lvl ? 1 : 0, highest_lvl ? *highest_lvl : 0, standalone ? 1 : 0,
                               max_res ? 1 : 0, max_res), "mg_fas_vcycle");
  sync_out(mg);
}

void mg_apply_op(mg_t& mg, int i_out) {                                           // :439-456
  sync_in(mg);
  gpu_check(mg, omg_apply_op(mg.gpu, i_out), "mg_apply_op");
  if (!mg.gpu_resident) mg_gpu_pull(mg, i_out, false);
}

// ------------------------------------------------------------------------------------------------
// m_diffusion: the three solvers differ only in the operator and in what lambda is scaled by
// ------------------------------------------------------------------------------------------------
static int diffusion_common(mg_t& mg, int op, double scale, int order, double max_res, double* res_out, const char* hint) {
  const int max_its = 10;                                                         // m_diffusion.f90:25
  if (order != 1 && order != 2) stop("diffusion_solve: order should be 1 or 2");
  mg.operator_type = op;
  mg_set_methods(mg);                                                             // :29-30
  // one push on entry, one pull on exit: the FMG cycle and the V-cycles below run device-resident
  if (!mg.gpu_resident) {
    need_gpu(mg);
    mg_gpu_push(mg, mg_iphi);
    mg_gpu_push(mg, mg_irhs);                                                     // order 1 multiplies it by f2 = 0: NaN/inf must behave as in the reference
    for (int e = 0; e < n_eps_vars(mg); ++e) mg_gpu_push(mg, mg_iveps + e);
  }
  if (order == 1) {                                                               // :33-35
    helmholtz_set_lambda(mg, 1 / scale);
    push_params(mg);
    gpu_check(mg, omg_diffusion_set_rhs(mg.gpu, -1 / scale, 0.0), "diffusion_solve");
  } else {                                                                        // :36-40
    helmholtz_set_lambda(mg, 0.0);
    push_params(mg);
    gpu_check(mg, omg_apply_op(mg.gpu, mg_irhs), "diffusion_solve");
    helmholtz_set_lambda(mg, 2 / scale);
    push_params(mg);
    gpu_check(mg, omg_diffusion_set_rhs(mg.gpu, -2 / scale, -1.0), "diffusion_solve");
  }
  double res = 0.0;
  gpu_check(mg, omg_fas_fmg(mg.gpu, 1, 1, &res), "diffusion_solve");              // :46
  int n;
  for (n = 1; n <= max_its; ++n) {                                                // :49-52
    if (res <= max_res) break;
    gpu_check(mg, omg_fas_vcycle(mg.gpu, 0, 0, 1, 1, &res), "diffusion_solve");
  }
  if (res_out) *res_out = res;
  sync_out(mg);
  if (n == max_its + 1) stop(std::string("diffusion_solve: no convergence (") + hint + ")");   // :54-58
  return n - 1;
}
int diffusion_solve(mg_t& mg, double dt, double diffusion_coeff, int order, double max_res, double* res) {   // :19-58
  return diffusion_common(mg, mg_helmholtz, dt * diffusion_coeff, order, max_res, res, "did you specify boundary conditions correctly?");
}
int diffusion_solve_vcoeff(mg_t& mg, double dt, int order, double max_res, double* res) {                    // :64-106
  return diffusion_common(mg, mg_vhelmholtz, dt, order, max_res, res, "boundary conditions? or is the variation in diffusion too large?");
}
int diffusion_solve_acoeff(mg_t& mg, double dt, int order, double max_res, double* res) {                    // :113-157
  return diffusion_common(mg, mg_ahelmholtz, dt, order, max_res, res, "boundary conditions? or is the variation in diffusion too large?");
}

void mg_fill_ghost_cells_lvl(mg_t& mg, int lvl, int iv) {                          // m_ghost_cells.f90:131-175
  push_params(mg);
  if (iv != mg_iphi && mg.bc[0][iv - 1].boundary_cond) stop("mg_fill_ghost_cells_lvl: boundary_cond callbacks are only supported for mg_iphi on the device");
  if (!mg.gpu_resident) transfer(mg, iv, true, false, lvl - 1, lvl);
  gpu_check(mg, omg_fill_ghost_cells_lvl(mg.gpu, lvl, iv), "mg_fill_ghost_cells_lvl");
  if (!mg.gpu_resident) transfer(mg, iv, false, true, lvl, lvl);
}
void mg_fill_ghost_cells(mg_t& mg, int iv) {                                      // :120-128
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) mg_fill_ghost_cells_lvl(mg, lvl, iv);
}

void mg_restrict_lvl(mg_t& mg, int iv, int lvl) {                                 // m_restrict.f90:83-114
  push_params(mg);
  if (!mg.gpu_resident) transfer(mg, iv, true, false, lvl - 1, lvl);
  gpu_check(mg, omg_restrict_lvl(mg.gpu, iv, lvl), "mg_restrict_lvl");
  if (!mg.gpu_resident) transfer(mg, iv, false, false, lvl - 1, lvl - 1);
}
void mg_restrict(mg_t& mg, int iv) {                                              // :72-80
  for (int lvl = mg.highest_lvl; lvl >= mg.lowest_lvl + 1; --lvl) mg_restrict_lvl(mg, iv, lvl);
}

void mg_prolong(mg_t& mg, int lvl, int iv, int iv_to, bool add) {                  // m_prolong.f90:51-85
  push_params(mg);
  if (!mg.gpu_resident) { transfer(mg, iv, true, false, lvl, lvl); transfer(mg, iv_to, true, false, lvl + 1, lvl + 1); }
  gpu_check(mg, omg_prolong(mg.gpu, lvl, iv, iv_to, add ? 1 : 0), "mg_prolong");
  if (!mg.gpu_resident) transfer(mg, iv_to, false, false, lvl + 1, lvl + 1);
}

// coupling_amrvac/mod_multigrid_coupling.t:133-171, 206-267 on the device tree (no host cc arrays in between)
int64_t mg_copy_size(const mg_t& mg, bool leaves_only, bool with_ghosts) {
  need_gpu(mg);
  return omg_copy_size(mg.gpu, mg_lvl_lo, mg_lvl_hi, leaves_only ? 1 : 0, with_ghosts ? 1 : 0);
}
void mg_copy_to_tree(mg_t& mg, int iv, const double* buf, int64_t n_doubles, bool leaves_only, bool with_ghosts) {
  if (!mg.is_allocated) stop("mg_copy_to_tree: tree not allocated yet");                  // :144-145
  push_params(mg);
  gpu_check(mg, omg_copy_to_tree(mg.gpu, iv, mg_lvl_lo, mg_lvl_hi, leaves_only ? 1 : 0, with_ghosts ? 1 : 0, buf, n_doubles), "mg_copy_to_tree");
}
void mg_copy_from_tree(mg_t& mg, int iv, double* buf, int64_t n_doubles, bool leaves_only, bool with_ghosts) {
  if (!mg.is_allocated) stop("mg_copy_from_tree: tree not allocated yet");                // :215-216
  push_params(mg);
  gpu_check(mg, omg_copy_from_tree(mg.gpu, iv, mg_lvl_lo, mg_lvl_hi, leaves_only ? 1 : 0, with_ghosts ? 1 : 0, buf, n_doubles), "mg_copy_from_tree");
}

}  // namespace omg

#include "m_free_space.inc"

// =====================================================================================================
// flat C wrapper for the Python harness (ctypes): one opaque pointer per mg_t
// =====================================================================================================
using namespace omg;
namespace {
struct Wrapped { mg_t mg; std::string err; };
typedef void (*py_bc_fn)(const double* r_min, const double* dr, int lvl, int nc, int iv, int nb, int* bc_type, double* bc, void* user);
struct PyBc { py_bc_fn fn; void* user; };
std::map<std::pair<void*, int>, PyBc>& py_bcs() { static std::map<std::pair<void*, int>, PyBc> m; return m; }
void bc_trampoline(const mg_box_t& box, int nc, int iv, int nb, int* bc_type, double* bc, void* user) {
  const PyBc* p = (const PyBc*)user;
  p->fn(box.r_min, box.dr, box.lvl, nc, iv, nb, bc_type, bc, p->user);
}
}  // namespace
#define H_TRY(w, ...) try { __VA_ARGS__; return 0; } catch (const std::exception& e) { (w)->err = e.what(); return 1; }

extern "C" {
void* omgh_create(int device) { Wrapped* w = new Wrapped(); w->mg.device = device; return w; }
void omgh_destroy(void* h) {
  Wrapped* w = (Wrapped*)h;
  for (int nb = 0; nb < 6; ++nb) py_bcs().erase({h, nb});
  delete w;
}
const char* omgh_last_error(void* h) { return ((Wrapped*)h)->err.c_str(); }
void* omgh_gpu_handle(void* h) { return ((Wrapped*)h)->mg.gpu; }

int omgh_set_int(void* h, const char* key, int v) {
  Wrapped* w = (Wrapped*)h; mg_t& mg = w->mg; const std::string k(key);
  H_TRY(w, {
    if (k == "operator_type") mg.operator_type = v; else if (k == "smoother_type") mg.smoother_type = v;
    else if (k == "geometry_type") mg.geometry_type = v;
    else if (k == "n_extra_vars") mg.n_extra_vars = v; else if (k == "n_cycle_down") mg.n_cycle_down = v;
    else if (k == "n_cycle_up") mg.n_cycle_up = v; else if (k == "max_coarse_cycles") mg.max_coarse_cycles = v;
    else if (k == "highest_lvl") mg.highest_lvl = v; else if (k == "gpu_resident") mg.gpu_resident = v != 0;
    else if (k == "coarsest_grid") { mg.coarsest_grid[0] = mg.coarsest_grid[1] = mg.coarsest_grid[2] = v; }
    else if (k == "subtract_mean") mg.subtract_mean = v != 0;
    else if (k == "bc_dirty") mg.bc_dirty = v != 0;
    else if (k == "host_cc") mg.host_cc = v != 0;
    else if (k == "coarse_collapse_boxes") mg.coarse_collapse_boxes = v;
    else throw mg_error("unknown int key " + k);
  });
}
int omgh_get_int(void* h, const char* key) {
  mg_t& mg = ((Wrapped*)h)->mg; const std::string k(key);
  if (k == "n_boxes") return mg.n_boxes; if (k == "lowest_lvl") return mg.lowest_lvl; if (k == "highest_lvl") return mg.highest_lvl;
  if (k == "first_normal_lvl") return mg.first_normal_lvl; if (k == "box_size") return mg.box_size;
  if (k == "n_extra_vars") return mg.n_extra_vars; if (k == "n_cpu") return mg.n_cpu; if (k == "my_rank") return mg.my_rank;
  if (k == "subtract_mean") return mg.subtract_mean; if (k == "n_smoother_substeps") return mg.n_smoother_substeps;
  if (k == "phi_bc_data_stored") return mg.phi_bc_data_stored;
  if (k == "gpu_resident") return mg.gpu_resident;
  return -999999;
}
int omgh_set_real(void* h, const char* key, double v) {
  Wrapped* w = (Wrapped*)h; mg_t& mg = w->mg; const std::string k(key);
  H_TRY(w, {
    if (k == "residual_coarse_abs") mg.residual_coarse_abs = v; else if (k == "residual_coarse_rel") mg.residual_coarse_rel = v;
    else if (k == "lambda") helmholtz_set_lambda(mg, v);
    else throw mg_error("unknown real key " + k);
  });
}
long long omgh_number_of_unknowns(void* h) { return mg_number_of_unknowns(((Wrapped*)h)->mg); }
int omgh_box_size_lvl(void* h, int lvl) { return ((Wrapped*)h)->mg.box_size_lvl[lvl - mg_lvl_lo]; }
void omgh_dr_lvl(void* h, int lvl, double* out) { for (int d = 0; d < 3; ++d) out[d] = ((Wrapped*)h)->mg.dr[lvl - mg_lvl_lo][d]; }
int omgh_set_bc(void* h, int nb, int iv, int bc_type, double bc_value) {
  mg_t& mg = ((Wrapped*)h)->mg; mg_bc_t& c = mg.bc[nb - 1][iv - 1];
  c.bc_type = bc_type; c.bc_value = bc_value; c.boundary_cond = nullptr; mg.bc_dirty = true; return 0;
}
int omgh_set_bc_callback(void* h, int nb, int iv, py_bc_fn fn, void* user) {
  mg_t& mg = ((Wrapped*)h)->mg;
  if (iv != mg_iphi) return 1;
  PyBc& slot = py_bcs()[{h, nb - 1}];
  slot = PyBc{fn, user};
  mg.bc[nb - 1][iv - 1].boundary_cond = bc_trampoline; mg.bc[nb - 1][iv - 1].user = &slot; mg.bc_dirty = true;
  return 0;
}
int omgh_comm_init(void* h, int my_rank, int n_cpu) { mg_comm_init(((Wrapped*)h)->mg, my_rank, n_cpu); return 0; }
int omgh_set_methods(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_set_methods(w->mg)); }
int omgh_build_rectangle(void* h, const int* domain_size, int box_size, const double* dx, const double* r_min, const int* periodic, int n_finer) {
  Wrapped* w = (Wrapped*)h; const bool per[3] = {periodic[0] != 0, periodic[1] != 0, periodic[2] != 0};
  H_TRY(w, mg_build_rectangle(w->mg, domain_size, box_size, dx, r_min, per, n_finer));
}
int omgh_add_children(void* h, int id) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_add_children(w->mg, id)); }
int omgh_set_leaves_parents(void* h, int lvl) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_set_leaves_parents(w->mg.boxes, w->mg.L(lvl))); }
int omgh_set_next_level_ids(void* h, int lvl) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_set_next_level_ids(w->mg, lvl)); }
int omgh_set_neighbors_lvl(void* h, int lvl) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_set_neighbors_lvl(w->mg, lvl)); }
int omgh_set_refinement_boundaries(void* h, int lvl) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_set_refinement_boundaries(w->mg.boxes, w->mg.L(lvl))); }
int omgh_load_balance(void* h, int kind) {
  Wrapped* w = (Wrapped*)h;
  H_TRY(w, { if (w->mg.n_cpu < 1) mg_comm_init(w->mg); if (kind == 0) mg_load_balance_simple(w->mg); else if (kind == 1) mg_load_balance(w->mg); else mg_load_balance_parents(w->mg); });
}
int omgh_allocate_storage(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_allocate_storage(w->mg)); }
int omgh_deallocate_storage(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_deallocate_storage(w->mg)); }
int omgh_lvl_list(void* h, int lvl, int which, int* out) {
  mg_lvl_t& L = ((Wrapped*)h)->mg.L(lvl);
  const std::vector<int>* v[8] = {&L.ids, &L.leaves, &L.parents, &L.ref_bnds, &L.my_ids, &L.my_leaves, &L.my_parents, &L.my_ref_bnds};
  if (out) std::copy(v[which]->begin(), v[which]->end(), out);
  return (int)v[which]->size();
}
void omgh_tree_export(void* h, int* out) {
  mg_t& mg = ((Wrapped*)h)->mg;
  for (int id = 1; id <= mg.n_boxes; ++id) {
    const mg_box_t& b = mg.boxes[id]; int* o = out + (size_t)(id - 1) * 20; int m = 0;
    o[m++] = b.rank; o[m++] = b.lvl; for (int d = 0; d < 3; ++d) o[m++] = b.ix[d]; o[m++] = b.parent;
    for (int c = 0; c < 8; ++c) o[m++] = b.children[c]; for (int f = 0; f < 6; ++f) o[m++] = b.neighbors[f];
  }
}
void omgh_box_rmin(void* h, int id, double* out) { for (int d = 0; d < 3; ++d) out[d] = ((Wrapped*)h)->mg.boxes[id].r_min[d]; }
void omgh_set_box_rank(void* h, int id, int rank) { ((Wrapped*)h)->mg.boxes[id].rank = rank; }
double* omgh_box_cc(void* h, int id) { mg_box_t& b = ((Wrapped*)h)->mg.boxes[id]; return b.cc.empty() ? nullptr : b.cc.data(); }

int omgh_fas_vcycle(void* h, int has_highest, int highest_lvl, int want_res, double* max_res, int standalone) {
  Wrapped* w = (Wrapped*)h;
  H_TRY(w, mg_fas_vcycle(w->mg, has_highest ? &highest_lvl : nullptr, want_res ? max_res : nullptr, standalone != 0));
}
int omgh_fas_fmg(void* h, int have_guess, int want_res, double* max_res) {
  Wrapped* w = (Wrapped*)h; H_TRY(w, mg_fas_fmg(w->mg, have_guess != 0, want_res ? max_res : nullptr));
}
int omgh_apply_op(void* h, int i_out) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_apply_op(w->mg, i_out)); }
int omgh_diffusion_solve(void* h, int kind, double dt, double coeff, int order, double max_res, double* res, int* n_vcycles) {
  Wrapped* w = (Wrapped*)h;
  H_TRY(w, { const int n = kind == 0 ? diffusion_solve(w->mg, dt, coeff, order, max_res, res)
                         : kind == 1 ? diffusion_solve_vcoeff(w->mg, dt, order, max_res, res)
                                     : diffusion_solve_acoeff(w->mg, dt, order, max_res, res);
             if (n_vcycles) *n_vcycles = n; });
}
int omgh_poisson_free_3d(void* h, int new_rhs, double max_fft_frac, int fmgcycle, int want_res, double* max_res) {
  Wrapped* w = (Wrapped*)h; H_TRY(w, mg_poisson_free_3d(w->mg, new_rhs != 0, max_fft_frac, fmgcycle != 0, want_res ? max_res : nullptr));
}
int omgh_free_space_fft_lvl(void* h) { Wrapped* w = (Wrapped*)h; return w->mg.free_bc_initialized ? w->mg.free_bc_fft_lvl : -999; }
int omgh_gpu_update_tree(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_gpu_update_tree(w->mg)); }
int omgh_fill_ghost_cells_lvl(void* h, int lvl, int iv) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_fill_ghost_cells_lvl(w->mg, lvl, iv)); }
int omgh_fill_ghost_cells(void* h, int iv) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_fill_ghost_cells(w->mg, iv)); }
int omgh_phi_bc_store(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_phi_bc_store(w->mg)); }
int omgh_restrict_lvl(void* h, int iv, int lvl) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_restrict_lvl(w->mg, iv, lvl)); }
int omgh_restrict(void* h, int iv) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_restrict(w->mg, iv)); }
int omgh_prolong(void* h, int lvl, int iv, int iv_to, int add) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_prolong(w->mg, lvl, iv, iv_to, add != 0)); }
int omgh_gpu_push(void* h, int iv) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_gpu_push(w->mg, iv)); }
int omgh_gpu_pull(void* h, int iv, int with_ghosts) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_gpu_pull(w->mg, iv, with_ghosts != 0)); }
int omgh_gpu_push_params(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_gpu_push_params(w->mg)); }
int omgh_gpu_push_bc(void* h) { Wrapped* w = (Wrapped*)h; H_TRY(w, mg_gpu_push_bc(w->mg)); }
}  // extern "C"
