// omg_plan.cpp -- static exchange plans (see omg_plan.h).  Pure host code: also used without a GPU
// by omg_exchange_plan() so that the partition logic can be tested on CPU-only ranks.
#include "omg_plan.h"

#include <algorithm>
#include <map>

namespace omgp {

namespace {
inline int rank_of(const omg_tree_desc& t, int id) { return t.box_rank[id - 1]; }
inline int nb_of(const omg_tree_desc& t, int id, int f) { return t.box_neighbors[(size_t)(id - 1) * 6 + f]; }
inline int child_number(const omg_tree_desc& t, int id) {
  // mg_ix_to_ichild, src/m_data_structures.f90:440-453 (1-based) -> 0-based
  const int32_t* ix = t.box_ix + (size_t)(id - 1) * 3;
  return 7 - 4 * (ix[2] & 1) - 2 * (ix[1] & 1) - (ix[0] & 1);
}
struct Lists { std::map<int, PeerPlan> by_peer; };
void add(Lists& L, int peer, bool send, const Item& it) {
  PeerPlan& p = L.by_peer[peer];
  p.peer = peer;
  (send ? p.send : p.recv).push_back(it);
}
}  // namespace

LevelPlan build_plan(const omg_tree_desc& t, int me, int ph, int lvl) {
  Lists L;
  LevelPlan out;
  if (lvl < t.lowest_lvl || lvl > t.highest_lvl || t.n_cpu <= 1) return out;
  const int a = t.lvl_ptr[lvl - OMG_LVL_LO], b = t.lvl_ptr[lvl - OMG_LVL_LO + 1];
  for (int q = a; q < b; ++q) {
    const int id = t.lvl_ids[q];
    const int r = rank_of(t, id);
    if (ph == PH_GHOST) {
      for (int f = 0; f < 6; ++f) {
        const int nb = nb_of(t, id, f);
        if (nb > 0) {
          // same-level face: receiver id gets the interior layer of nb (buffer_for_nb / fill_buffered_nb)
          const int rs = rank_of(t, nb);
          if (rs == r) continue;
          const Item it{(int64_t)6 * id + f, id, nb, f, IT_SAME};
          if (r == me) add(L, rs, false, it);
          if (rs == me) add(L, r, true, it);
        } else if (nb == 0) {
          // refinement boundary: the coarse neighbour of my parent interpolates the face
          // (buffer_for_fine_nb, src/m_ghost_cells.f90:385-422)
          const int p = t.box_parent[id - 1];
          if (p <= 0) continue;
          const int cnb = nb_of(t, p, f);
          if (cnb <= 0) continue;
          const int rs = rank_of(t, cnb);
          if (rs == r) continue;
          const Item it{(int64_t)6 * id + f, id, cnb, f, IT_RB};
          if (r == me) add(L, rs, false, it);
          if (rs == me) add(L, r, true, it);
        }
      }
    } else {
      const int p = t.box_parent[id - 1];
      if (p <= 0) continue;
      const int rp = rank_of(t, p);
      if (rp == r) continue;
      if (ph == PH_RESTRICT) {      // child id (sender r) -> parent p (receiver rp)
        const int c = child_number(t, id);
        const Item it{(int64_t)8 * p + c, p, id, c, IT_SAME};
        if (r == me) add(L, rp, true, it);
        if (rp == me) add(L, r, false, it);
      } else {                      // parent p (sender rp) -> child id (receiver r)
        const Item it{(int64_t)id, id, p, child_number(t, id), IT_SAME};
        if (rp == me) add(L, r, true, it);
        if (r == me) add(L, rp, false, it);
      }
    }
  }
  // message order: same-level faces first, then refinement-boundary faces, each by ascending key
  auto cmp = [](const Item& x, const Item& y) { return x.kind != y.kind ? x.kind < y.kind : x.key < y.key; };
  for (auto& kv : L.by_peer) {
    std::sort(kv.second.send.begin(), kv.second.send.end(), cmp);
    std::sort(kv.second.recv.begin(), kv.second.recv.end(), cmp);
    out.peers.push_back(std::move(kv.second));
  }
  return out;
}

}  // namespace omgp
