// omg_fs_kernel.cpp -- host-side construction of the free-space kernel (include/omg_gpu.h: omg_free_space_setup): the
// kernel of 1/r in separable form -- what the reference's createKernel computes before its FFT
// (poisson_3d_fft/build_kernel.f90:884-1050, from BigDFT): 1/r ~ a sum of 89 Gaussians (gequad), each integrated
// against the order-8 interpolating scaling function per direction.  Pure host code; hands the result to
// omg_free_space_kernel.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "../../include/omg_gpu.h"
#include "gequad_table.inc"

namespace {

// Deslauriers-Dubuc interpolation of order 8: the value midway between grid points from the 8 nearest ones.  These are
// the odd taps of the reference's filter ch (poisson_3d_fft/lazy_8.inc); the even taps are the identity.
const double kIsf8Odd[8] = {-5.0 / 2048.0, 49.0 / 2048.0, -245.0 / 2048.0, 1225.0 / 2048.0,
                            1225.0 / 2048.0, -245.0 / 2048.0, 49.0 / 2048.0, -5.0 / 2048.0};   // offsets -7, -5, ..., +7 (in half steps)

// the order-8 interpolating scaling function sampled at n + 1 points of [-7, 9] (n = 16 * 2^q): cascade from the unit
// sample at 0, each step halving the spacing (scaling_function + back_trans_8, scaling_function.f90:18-96, 328-370;
// the periodic wrap of the reference never carries a non-zero value, the support is [-7, 7])
std::vector<double> isf8_samples(int n) {
  std::vector<double> cur(16, 0.0);
  cur[7] = 1.0;                                         // x = -7 + i
  while ((int)cur.size() < n) {
    const int m = (int)cur.size();
    std::vector<double> nxt((size_t)2 * m, 0.0);
    for (int i = 0; i < m; ++i) {
      nxt[2 * i] = nxt[2 * i] + 1.0 * cur[i];           // ch(0) = 1
      double odd = 0.0;                                  // y(2i+1) = sum_j ch(2j+1) x(i-j), j = -5 .. 4 (taps +-9 are zero)
      for (int j = -5; j <= 4; ++j) {
        const int t = 2 * j + 1;                         // tap index -9 .. 9
        const double c = (t >= -7 && t <= 7) ? kIsf8Odd[(t + 7) / 2] : 0.0;
        int ind = i - j;
        ind = ((ind % m) + m) % m;
        odd = odd + c * cur[ind];
      }
      nxt[2 * i + 1] = odd;
    }
    cur.swap(nxt);
  }
  cur.push_back(0.0);                                   // the reference's arrays run 0 .. n
  return cur;
}

// one refinement step of the kernel of a Gaussian 4x narrower (scf_recursion_8, scaling_function.f90:443-479):
// k'(i) = 1/2 sum_j ch(j) k(2i - j), symmetric, zero beyond the range
void refine_kernel(std::vector<double>& k, int n_range, int steps) {
  std::vector<double> old;
  auto at = [&](const std::vector<double>& v, int i) { i = std::abs(i); return i > n_range ? 0.0 : v[i]; };
  for (int s = 0; s < steps; ++s) {
    old = k;
    std::fill(k.begin(), k.end(), 0.0);
    for (int i = 0; i <= n_range; ++i) {
      double tot = 0.0;
      for (int j = -10; j <= 10; ++j) {
        const double c = j == 0 ? 1.0 : ((j & 1) && j >= -7 && j <= 7 ? kIsf8Odd[(j + 7) / 2] : 0.0);
        tot = tot + c * at(old, 2 * i - j);
      }
      if (tot == 0.0) break;
      k[i] = 0.5 * tot;
    }
  }
}

struct FreeKernel { int n_range = 0; std::vector<double> w, k[3]; };   // k[d][g * (n_range + 1) + i], i >= 0 (symmetric)

// Free_Kernel up to its FFT (build_kernel.f90:884-1050): for every Gaussian exp(-p r^2) of the quadrature, the integrals
// k(i) = int phi(x) exp(-p h^2 (x - i)^2) dx per direction -- computed for a Gaussian wide enough for the 1024-point
// rule and refined to the requested width -- so that K(i, j, k) = sum_g w_g k_g(i) k_g(j) k_g(k)
FreeKernel build_free_kernel(int n_max, const double h[3], const double len[3]) {
  const int n_gauss = 89, n_scf = 2 * 8 * 64;
  const std::vector<double> y = isf8_samples(n_scf);
  std::vector<double> x(n_scf + 1);
  for (int i = 0; i <= n_scf; ++i) x[i] = (double)(i * 16) / (double)n_scf - (0.5 * 16.0 - 1.0);
  const double dx = 16.0 / (double)n_scf;
  FreeKernel K;
  K.n_range = std::max(n_max, 16);
  const int nr1 = K.n_range + 1;
  const double l2 = len[0] * len[0] + len[1] * len[1] + len[2] * len[2];
  K.w.resize(n_gauss);
  for (int d = 0; d < 3; ++d) K.k[d].assign((size_t)n_gauss * nr1, 0.0);
  const bool cubic = (h[0] == h[1] && h[1] == h[2]);
  for (int g = n_gauss - 1; g >= 0; --g) {
    const double p = kGequadP[g] * (1.0 / l2);
    K.w[g] = kGequadW[g] * (1.0 / std::sqrt(l2));
    for (int d = 0; d < (cubic ? 1 : 3); ++d) {
      const double p_cell = 1.0 / (h[d] * h[d]);
      int steps = (int)std::lround((std::log(p) - std::log(p_cell)) / std::log(4.0));
      if (steps < 0) steps = 0;
      const double p0 = steps > 0 ? p / std::pow(4.0, steps) : p;
      std::vector<double> k(nr1, 0.0);
      for (int c = 0; c <= K.n_range; ++c) {
        double s = 0.0;
        for (int i = 0; i <= n_scf; ++i) {
          const double t = x[i] - (double)c;
          s = s + y[i] * std::exp(cubic ? -p0 * (t * t * (h[d] * h[d])) : -p0 * t * t * (h[d] * h[d]));
        }
        k[c] = s * dx;
        if (std::fabs(s) < 1e-18) break;
      }
      refine_kernel(k, K.n_range, steps);
      std::copy(k.begin(), k.end(), K.k[d].begin() + (size_t)g * nr1);
    }
    if (cubic) for (int d = 1; d < 3; ++d) std::copy(K.k[0].begin() + (size_t)g * nr1, K.k[0].begin() + (size_t)(g + 1) * nr1, K.k[d].begin() + (size_t)g * nr1);
  }
  return K;
}

}  // namespace

extern "C" int omg_free_space_setup(omg_handle* h, int fft_lvl, const double r_min[3], const int n_kernel[3], const double dr[3]) {
  if (!h || !r_min || !n_kernel || !dr) return 1;
  int nmax = 0; double len[3];
  for (int d = 0; d < 3; ++d) { nmax = std::max(nmax, n_kernel[d]); len[d] = dr[d] * (double)n_kernel[d]; }
  const FreeKernel K = build_free_kernel(nmax, dr, len);
  return omg_free_space_kernel(h, fft_lvl, r_min, (int)K.w.size(), K.n_range, K.w.data(), K.k[0].data(), K.k[1].data(), K.k[2].data());
}
