// omg_plan.h -- static exchange plans of the three communication phases (host only, no CUDA).
//
// The reference packs its send buffers at run time and sorts them by a destination key
// (src/m_ghost_cells.f90:348-422 key 6*nb_id + rev(nb); src/m_restrict.f90:116-163 key 8*p_id + ichild;
// src/m_prolong.f90:91-121 key c_id; sort: src/m_communication.f90:68-88).  The tree is replicated on
// every rank, so each rank can enumerate -- once, at setup -- both the items it sends to a peer and the
// items it receives from it, in the same ascending-key order.  The cycle then runs pack kernels over
// these lists and one ncclSend/ncclRecv pair per peer and phase; there is no run-time sort.
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/omg_gpu.h"

namespace omgp {

enum Phase : int { PH_GHOST = 0, PH_RESTRICT = 1, PH_PROLONG = 2 };
enum ItemKind : int { IT_SAME = 0, IT_RB = 1 };

// One message item.  `key` orders the items of a (phase, level, peer) message on both sides.
//  ghost   : box = receiving box (global id), face 0..5 of the RECEIVER, src = sending box
//            (IT_SAME: same-level neighbour; IT_RB: coarse neighbour of the receiver's parent)
//  restrict: box = parent (receiver side), src = child (sender side), face = child number 0..7
//  prolong : box = child (receiver side), src = parent (sender side)
struct Item { int64_t key; int box, src, face, kind; };

struct PeerPlan { int peer = -1; std::vector<Item> send, recv; };
struct LevelPlan { std::vector<PeerPlan> peers; };   // ascending peer rank; only peers with traffic

// Plan of phase ph for level lvl as seen from rank `me`:
//  PH_GHOST    ghost cells of level lvl (same-level faces on lvl, coarse->fine faces from lvl-1)
//  PH_RESTRICT children on lvl -> parents on lvl-1
//  PH_PROLONG  parents on lvl-1 -> children on lvl
LevelPlan build_plan(const omg_tree_desc& t, int me, int ph, int lvl);

}  // namespace omgp
