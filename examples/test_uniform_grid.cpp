// test_uniform_grid.cpp -- the reference's tests/test_uniform_grid.f90 written against the C++ host mirror
// (octree-mg_b200/host/m_octree_mg.hpp): same set-up calls, same manufactured solution, same output lines.
// What a C++ host of the drop-in boundary looks like; the Fortran equivalent only swaps `use m_octree_mg`'s cycle
// routines for m_gpu_multigrid.  There is no CPU fallback: without a CUDA device
// mg_allocate_storage stops with the library's message.
//
//   ./test_uniform_grid box_size nx ny nz [n_its] [FMG? t/f] [smoother: gs|gsrb]
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../octree-mg_b200/host/m_octree_mg.hpp"

using namespace omg;

namespace {
const int n_modes[3] = {5, 5, 5};                       // tests/test_uniform_grid.f90:17
const double pi = std::acos(-1.0);
int i_eps, i_sol;

double solution(const double r[3]) {                     // :131-134
  double p = 1.0;
  for (int d = 0; d < 3; ++d) p *= std::sin(2 * pi * n_modes[d] * r[d]);
  return p;
}

inline size_t cix(int n2, int i, int j, int k, int iv) { return (((size_t)(iv - 1) * n2 + k) * n2 + j) * n2 + i; }

void set_solution(mg_t& mg) {                            // :136-152
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo], n2 = nc + 2;
    for (int id : mg.L(lvl).my_ids) {
      mg_box_t& box = mg.boxes[id];
      for (int k = 0; k <= nc + 1; ++k) for (int j = 0; j <= nc + 1; ++j) for (int i = 0; i <= nc + 1; ++i) {
        const int ijk[3] = {i, j, k};
        double r[3];
        for (int d = 0; d < 3; ++d) r[d] = box.r_min[d] + (ijk[d] - 0.5) * mg.dr[lvl - mg_lvl_lo][d];
        box.cc[cix(n2, i, j, k, i_sol)] = solution(r);
        box.cc[cix(n2, i, j, k, i_eps)] = std::fmax(1.0, 1.0 + r[0]);
      }
    }
  }
}

// compute_rhs_and_reset :154-169: rhs = box_lpl(solution incl. its analytic ghost cells) evaluated on the host, as the
// reference does through mg%box_op (src/m_laplacian.f90:183-189), then phi = 0
void compute_rhs_and_reset(mg_t& mg) {
  for (int lvl = mg.lowest_lvl; lvl <= mg.highest_lvl; ++lvl) {
    const int nc = mg.box_size_lvl[lvl - mg_lvl_lo], n2 = nc + 2;
    double idr2[3];
    for (int d = 0; d < 3; ++d) idr2[d] = 1.0 / (mg.dr[lvl - mg_lvl_lo][d] * mg.dr[lvl - mg_lvl_lo][d]);
    for (int id : mg.L(lvl).my_ids) {
      double* cc = mg.boxes[id].cc.data();
      for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
        const double u0 = cc[cix(n2, i, j, k, i_sol)];
        cc[cix(n2, i, j, k, mg_irhs)] =
            idr2[0] * (cc[cix(n2, i - 1, j, k, i_sol)] + cc[cix(n2, i + 1, j, k, i_sol)] - 2 * u0) +
            idr2[1] * (cc[cix(n2, i, j - 1, k, i_sol)] + cc[cix(n2, i, j + 1, k, i_sol)] - 2 * u0) +
            idr2[2] * (cc[cix(n2, i, j, k - 1, i_sol)] + cc[cix(n2, i, j, k + 1, i_sol)] - 2 * u0);
      }
      for (int k = 0; k <= nc + 1; ++k) for (int j = 0; j <= nc + 1; ++j) for (int i = 0; i <= nc + 1; ++i)
        cc[cix(n2, i, j, k, mg_iphi)] = 0.0;
    }
  }
}

void print_error(mg_t& mg, int it) {                     // :171-202 (one rank: no mpi_reduce)
  double err = 0.0, res = 0.0;
  const int lvl = mg.highest_lvl, nc = mg.box_size_lvl[lvl - mg_lvl_lo], n2 = nc + 2;
  for (int id : mg.L(lvl).my_ids) {
    const double* cc = mg.boxes[id].cc.data();
    for (int k = 1; k <= nc; ++k) for (int j = 1; j <= nc; ++j) for (int i = 1; i <= nc; ++i) {
      err = std::fmax(err, std::fabs(cc[cix(n2, i, j, k, mg_iphi)] - cc[cix(n2, i, j, k, i_sol)]));
      res = std::fmax(res, std::fabs(cc[cix(n2, i, j, k, mg_ires)]));
    }
  }
  std::printf("%4d  max solution error%12.4E  max residual%12.4E\n", it, err, res);
}

// my_boundary_condition :204-220: Dirichlet values of the solution at the face centres
void my_boundary_condition(const mg_box_t& box, int nc, int /*iv*/, int nb, int* bc_type, double* bc, void* /*user*/) {
  std::vector<double> x(3 * (size_t)nc * nc);
  mg_get_face_coords(box, nb, nc, x.data());
  *bc_type = mg_bc_dirichlet;
  for (int t = 0; t < nc * nc; ++t) {
    const double r[3] = {x[t], x[(size_t)nc * nc + t], x[2 * (size_t)nc * nc + t]};
    bc[t] = solution(r);
  }
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 5 || argc > 8) {
    std::fprintf(stderr, "Usage: ./test_uniform_grid box_size domain_size(3) [n_its] [FMG?] [gs|gsrb]\n");
    return 2;
  }
  const int box_size = std::atoi(argv[1]);
  const int domain_size[3] = {std::atoi(argv[2]), std::atoi(argv[3]), std::atoi(argv[4])};
  const int n_its = argc > 5 ? std::atoi(argv[5]) : 10;
  const bool fmg_cycle = argc > 6 ? (argv[6][0] == 't' || argv[6][0] == 'T') : true;
  const bool gsrb = argc > 7 ? std::strcmp(argv[7], "gsrb") == 0 : false;
  double dr[3];
  const double r_min[3] = {0.0, 0.0, 0.0};
  const bool periodic[3] = {false, false, false};
  for (int d = 0; d < 3; ++d) dr[d] = 1.0 / domain_size[d];

  try {
    mg_t mg;
    mg.n_extra_vars = 2;
    i_eps = mg_num_vars + 1;
    i_sol = mg_num_vars + 2;
    mg.geometry_type = mg_cartesian;
    mg.operator_type = mg_laplacian;
    mg.smoother_type = gsrb ? mg_smoother_gsrb : mg_smoother_gs;     // the reference program uses mg_smoother_gs (:64)
    for (int n = 0; n < mg_num_neighbors; ++n) mg.bc[n][mg_iphi - 1].boundary_cond = my_boundary_condition;

    mg_set_methods(mg);
    mg_comm_init(mg);
    auto t0 = std::chrono::steady_clock::now();
    mg_build_rectangle(mg, domain_size, box_size, dr, r_min, periodic, 0);
    mg_load_balance(mg);
    auto t1 = std::chrono::steady_clock::now();
    mg_allocate_storage(mg);                              // host cc arrays + the device mirror (fails loudly without a GPU)
    auto t2 = std::chrono::steady_clock::now();
    std::printf(" mesh construction (s) %g\n allocate storage (s)  %g\n", std::chrono::duration<double>(t1 - t0).count(),
                std::chrono::duration<double>(t2 - t1).count());

    set_solution(mg);
    compute_rhs_and_reset(mg);
    print_error(mg, 0);

    t0 = std::chrono::steady_clock::now();
    for (int n = 1; n <= n_its; ++n) {
      if (fmg_cycle) mg_fas_fmg(mg, n > 1); else mg_fas_vcycle(mg);   // strict drop-in mode: cc is synced in and out
      print_error(mg, n);
    }
    t1 = std::chrono::steady_clock::now();
    const double dt = std::chrono::duration<double>(t1 - t0).count();
    std::printf(" cycle type        %s\n n_cpu             %d\n problem_size      %d %d %d\n box_size          %d\n"
                " n_iterations      %d\n time/iteration    %g\n total_time(s)     %g\n unknowns/microsec %g\n",
                fmg_cycle ? "FMG" : "V-cycle", mg.n_cpu, domain_size[0], domain_size[1], domain_size[2], box_size, n_its,
                dt / n_its, dt, 1e-6 * n_its * (double)mg_number_of_unknowns(mg) / dt);
    mg_deallocate_storage(mg);
  } catch (const mg_error& e) {
    std::fprintf(stderr, "ERROR STOP %s\n", e.what());    // the reference `error stop`s
    return 1;
  }
  return 0;
}
