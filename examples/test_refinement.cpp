// test_refinement.cpp -- the reference's tests/test_refinement.f90 written against the C++ host mirror: a base grid
// refined n_levels - 1 times around the domain centre (mg_add_children + the level bookkeeping calls of
// build_amr_tree, :194-247), manufactured solution, Dirichlet-0 boundaries, FMG or V-cycles, the error on the leaves
// printed after every cycle, the level table and the timing lines at the end.
//
//   ./test_refinement n_levels box_size nx ny nz [n_its] [FMG? t/f] [smoother: gs|gsrb]
//
// One difference, as in tests/problems.py: the right-hand side is the discrete operator of the solution evaluated by
// the library (mg_apply_op on the device, where phi's ghost cells at refinement boundaries are derived) instead of
// per-box calls of mg%box_op on host ghost cells (:147-163) -- the same numbers, there is no host operator here.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../octree-mg_b200/host/m_octree_mg.hpp"

using namespace omg;

namespace {
const int n_modes[3] = {5, 5, 5};                        // tests/test_refinement.f90:17
const double pi = std::acos(-1.0);
int i_sol;

inline size_t cix(int n2, int i, int j, int k, int iv) { return (((size_t)(iv - 1) * n2 + k) * n2 + j) * n2 + i; }

// build_amr_tree :194-247
void build_amr_tree(mg_t& mg, int n_amr_levels, const int lvl1_size[3], int box_size, const double dr[3], const bool periodic[3]) {
  const double r_min[3] = {0.0, 0.0, 0.0};
  double domain_len[3];
  int n_finer = 1000, per_lvl = 1;
  for (int d = 0; d < 3; ++d) { domain_len[d] = lvl1_size[d] * dr[d]; per_lvl *= lvl1_size[d] / box_size; }
  n_finer += n_amr_levels * per_lvl;                     // estimate of the boxes on refined levels
  mg_build_rectangle(mg, lvl1_size, box_size, dr, r_min, periodic, n_finer);

  for (int lvl = 1; lvl <= n_amr_levels - 1; ++lvl) {
    const std::vector<int> ids = mg.L(lvl).ids;          // copy: mg_add_children appends boxes
    for (int id : ids) {
      bool inside = true;                                // refine the centre region
      for (int d = 0; d < 3; ++d) {
        const double r0 = 0.5 * domain_len[d] - domain_len[d] * std::pow(0.5, lvl + 1);
        const double r1 = 0.5 * domain_len[d] + domain_len[d] * std::pow(0.5, lvl + 1);
        const double centre = mg.boxes[id].r_min[d] + 0.5 * box_size * mg.boxes[id].dr[d];
        inside = inside && centre >= r0 && centre <= r1;
      }
      if (inside) mg_add_children(mg, id);
    }
    mg_set_leaves_parents(mg.boxes, mg.L(lvl));
    mg_set_next_level_ids(mg, lvl);
    mg_set_neighbors_lvl(mg, lvl + 1);
  }
  mg_set_leaves_parents(mg.boxes, mg.L(n_amr_levels));
  mg.highest_lvl = n_amr_levels;
  for (int lvl = 1; lvl <= mg.highest_lvl; ++lvl) mg_set_refinement_boundaries(mg.boxes, mg.L(lvl));   // seen from the coarse side
  mg_load_balance(mg);
  mg_allocate_storage(mg);                               // host cc arrays + the device mirror (stops loudly without a GPU)
}

// set_solution :124-145 and compute_rhs_and_reset :147-163
void set_solution_and_rhs(mg_t& mg) {
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo], n2 = nc + 2;
    for (int id : mg.L(lvl).my_ids) {
      mg_box_t& box = mg.boxes[id];
      for (int k = 0; k <= nc + 1; ++k) for (int j = 0; j <= nc + 1; ++j) for (int i = 0; i <= nc + 1; ++i) {
        const int ijk[3] = {i, j, k};
        double sol = 1.0;
        for (int d = 0; d < 3; ++d) sol *= std::sin(2 * pi * n_modes[d] * (box.r_min[d] + (ijk[d] - 0.5) * mg.dr[lvl - mg_lvl_lo][d]));
        box.cc[cix(n2, i, j, k, i_sol)] = sol;
      }
    }
  }
  // parents carry the restricted solution and ghost cells follow from the library's own fill, as in the reference
  mg_restrict(mg, i_sol);
  mg_fill_ghost_cells(mg, i_sol);
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int n2 = mg.box_size_lvl[lvl - mg_lvl_lo] + 2;
    for (int id : mg.L(lvl).my_ids) {
      double* cc = mg.boxes[id].cc.data();
      std::memcpy(cc + cix(n2, 0, 0, 0, mg_iphi), cc + cix(n2, 0, 0, 0, i_sol), sizeof(double) * n2 * n2 * n2);
    }
  }
  mg_fill_ghost_cells(mg, mg_iphi);
  mg_apply_op(mg, mg_irhs);                              // rhs = L(solution) on every box of every level
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int n2 = mg.box_size_lvl[lvl - mg_lvl_lo] + 2;
    for (int id : mg.L(lvl).my_ids) {
      double* phi = mg.boxes[id].cc.data() + cix(n2, 0, 0, 0, mg_iphi);
      for (size_t q = 0; q < (size_t)n2 * n2 * n2; ++q) phi[q] = 0.0;
    }
  }
}

void print_error(mg_t& mg, int it) {                     // :165-192, leaves of levels 1..highest (one rank: no mpi_reduce)
  double err = 0.0;
  for (int lvl = 1; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo], n2 = nc + 2;
    for (int id : mg.L(lvl).my_leaves) {
      const double* cc = mg.boxes[id].cc.data();
      for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i)
        err = std::fmax(err, std::fabs(cc[cix(n2, i, j, k, mg_iphi)] - cc[cix(n2, i, j, k, i_sol)]));
    }
  }
  std::printf("%12d max err%24.16E\n", it, err);
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 6 || argc > 9) {
    std::fprintf(stderr, "Usage: ./test_refinement n_levels box_size domain_size(3) [n_its] [FMG?] [gs|gsrb]\n");
    return 2;
  }
  const int n_levels = std::atoi(argv[1]), box_size = std::atoi(argv[2]);
  const int domain_size[3] = {std::atoi(argv[3]), std::atoi(argv[4]), std::atoi(argv[5])};
  const int n_its = argc > 6 ? std::atoi(argv[6]) : 10;
  const bool fmg_cycle = argc > 7 ? (argv[7][0] == 't' || argv[7][0] == 'T') : true;
  const bool gsrb = argc > 8 ? std::strcmp(argv[8], "gsrb") == 0 : false;
  double dr[3];
  const bool periodic[3] = {false, false, false};
  for (int d = 0; d < 3; ++d) dr[d] = 1.0 / domain_size[d];

  try {
    mg_t mg;
    mg.n_extra_vars = 1;
    i_sol = mg_num_vars + 1;
    mg.geometry_type = mg_cartesian;
    mg.operator_type = mg_laplacian;
    mg.smoother_type = gsrb ? mg_smoother_gsrb : mg_smoother_gs;      // the reference program uses mg_smoother_gs (:64)
    for (int n = 0; n < mg_num_neighbors; ++n) {
      mg.bc[n][mg_iphi - 1].bc_type = mg_bc_dirichlet; mg.bc[n][mg_iphi - 1].bc_value = 0.0;
      mg.bc[n][i_sol - 1].bc_type = mg_bc_dirichlet; mg.bc[n][i_sol - 1].bc_value = 0.0;
    }
    mg_set_methods(mg);
    mg_comm_init(mg);
    build_amr_tree(mg, n_levels, domain_size, box_size, dr, periodic);
    set_solution_and_rhs(mg);
    print_error(mg, 0);

    const auto t0 = std::chrono::steady_clock::now();
    for (int n = 1; n <= n_its; ++n) {
      if (fmg_cycle) mg_fas_fmg(mg, n > 1); else mg_fas_vcycle(mg);   // strict drop-in mode: cc is synced in and out
      print_error(mg, n);
    }
    const auto t1 = std::chrono::steady_clock::now();
    const double dt = std::chrono::duration<double>(t1 - t0).count();

    for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl)
      std::printf(" lvl_%d: %zu boxes, %zu leaves, box size: %d\n", lvl, mg.L(lvl).ids.size(), mg.L(lvl).leaves.size(),
                  mg.box_size_lvl[lvl - mg_lvl_lo]);
    const double n_unknowns = (double)mg.n_boxes * box_size * box_size * box_size;
    std::printf(" cycle type        %s\n n_cpu             %d\n coarse grid       %d %d %d\n box_size          %d\n"
                " n_iterations      %d\n time/iteration    %g\n total_time(s)     %g\n unknowns/microsec %g\n",
                fmg_cycle ? "FMG" : "V-cycle", mg.n_cpu, domain_size[0], domain_size[1], domain_size[2], box_size, n_its,
                dt / n_its, dt, 1e-6 * n_its * n_unknowns / dt);
    mg_deallocate_storage(mg);
  } catch (const mg_error& e) {
    std::fprintf(stderr, "ERROR STOP %s\n", e.what());
    return 1;
  }
  return 0;
}
