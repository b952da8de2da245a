// omg_fs.cu -- see omg_fs.h.  The reference solves the free-space problem of the FFT level with the BigDFT Poisson
// solver vendored under poisson_3d_fft/ (PSolver, geocode 'F': psolver_main.f90:296-300): potential = hx hy hz * (K conv
// rho) with K the kernel of 1/r in a basis of interpolating scaling functions.  Here the same linear convolution is one
// zero-padded real-to-complex cuFFT, a pointwise product with the transformed kernel and the inverse transform; the
// kernel's one-dimensional factors come from the host (host/m_free_space.cpp).
#include "omg_fs.h"

#include <cstdlib>
#include <mutex>

#include <cufft.h>   // types only: the library is bound at run time (one copy per process, as for NCCL)
#include <dlfcn.h>

namespace omgfs {
namespace {

struct CufftApi {
  cufftResult (*Plan3d)(cufftHandle*, int, int, int, cufftType) = nullptr;
  cufftResult (*ExecD2Z)(cufftHandle, cufftDoubleReal*, cufftDoubleComplex*) = nullptr;
  cufftResult (*ExecZ2D)(cufftHandle, cufftDoubleComplex*, cufftDoubleReal*) = nullptr;
  cufftResult (*SetStream)(cufftHandle, cudaStream_t) = nullptr;
  cufftResult (*Destroy)(cufftHandle) = nullptr;
  std::string err;
};
const CufftApi& cufft_api() {
  static CufftApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* env = getenv("OMG_CUFFT_LIB");
    const char* names[] = {env && *env ? env : "libcufft.so.11", "libcufft.so.11", "libcufft.so", "/usr/local/cuda/lib64/libcufft.so.11",
                           "/usr/local/cuda/lib64/libcufft.so"};
    void* lib = nullptr;
    for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_LOCAL); if (lib) break; }
    if (!lib) { api.err = std::string("cannot load cuFFT: ") + dlerror(); return; }
    auto sym = [&](const char* n) { void* p = dlsym(lib, n); if (!p && api.err.empty()) api.err = std::string("cuFFT symbol missing: ") + n; return p; };
    api.Plan3d = (decltype(api.Plan3d))sym("cufftPlan3d");
    api.ExecD2Z = (decltype(api.ExecD2Z))sym("cufftExecD2Z");
    api.ExecZ2D = (decltype(api.ExecZ2D))sym("cufftExecZ2D");
    api.SetStream = (decltype(api.SetStream))sym("cufftSetStream");
    api.Destroy = (decltype(api.Destroy))sym("cufftDestroy");
  });
  return api;
}

constexpr int kT = 256;

// real-space kernel on the padded grid, wrap-around placement: offset d lives at index d and at M - d
__global__ void k_build_kernel(double* __restrict__ K, int M1, int M2, int M3, int n1, int n2, int n3, int n_gauss, int nr1,
                               const double* __restrict__ w, const double* __restrict__ kx, const double* __restrict__ ky,
                               const double* __restrict__ kz) {
  const size_t N = (size_t)M1 * M2 * M3;
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < N; q += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(q % M1), j = (int)((q / M1) % M2), k = (int)(q / ((size_t)M1 * M2));
    const int di = min(i, M1 - i), dj = min(j, M2 - j), dk = min(k, M3 - k);
    double v = 0.0;
    if (di < n1 && dj < n2 && dk < n3 && di < nr1 && dj < nr1 && dk < nr1)
      for (int g = n_gauss - 1; g >= 0; --g) v = v + w[g] * kx[(size_t)g * nr1 + di] * ky[(size_t)g * nr1 + dj] * kz[(size_t)g * nr1 + dk];
    K[q] = v;
  }
}
__global__ void k_pad(double* __restrict__ pad, const double* __restrict__ grid, int M1, int M2, int M3, int n1, int n2, int n3) {
  const size_t N = (size_t)M1 * M2 * M3;
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < N; q += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(q % M1), j = (int)((q / M1) % M2), k = (int)(q / ((size_t)M1 * M2));
    pad[q] = (i < n1 && j < n2 && k < n3) ? grid[((size_t)k * n2 + j) * n1 + i] : 0.0;
  }
}
__global__ void k_mul_spec(cufftDoubleComplex* __restrict__ a, const cufftDoubleComplex* __restrict__ b, size_t n) {
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < n; q += (size_t)gridDim.x * blockDim.x) {
    const cufftDoubleComplex x = a[q], y = b[q];
    a[q].x = x.x * y.x - x.y * y.y;
    a[q].y = x.x * y.y + x.y * y.x;
  }
}
__global__ void k_crop(double* __restrict__ grid, const double* __restrict__ pad, int M1, int M2, int n1, int n2, int n3, double scal) {
  const size_t N = (size_t)n1 * n2 * n3;
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < N; q += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(q % n1), j = (int)((q / n1) % n2), k = (int)(q / ((size_t)n1 * n2));
    grid[q] = scal * pad[((size_t)k * M2 + j) * M1 + i];
  }
}
inline int grid_for(size_t n) { size_t g = (n + kT - 1) / kT; return (int)(g > 148 * 32 ? 148 * 32 : (g == 0 ? 1 : g)); }

// rhs of the FFT level -> dense grid (m_free_space.f90:141-150); one CTA per box
__global__ void __launch_bounds__(kT) k_gather_rhs(const __grid_constant__ LevelDev L, const int* __restrict__ origin, double rhs_fac,
                                                   double* __restrict__ grid, int n1, int n2) {
  const int b = blockIdx.x;
  const double* v = L.var[1] + (size_t)b * L.bs;
  const int ox = origin[3 * b] + 1, oy = origin[3 * b + 1] + 1, oz = origin[3 * b + 2] + 1;   // + the grid's ghost layer
  for (int idx = threadIdx.x; idx < L.bs; idx += blockDim.x) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    grid[((size_t)(oz + k) * n2 + (oy + j)) * n1 + (ox + i)] = rhs_fac * v[idx];
  }
}
// the solution as initial guess on the FFT level (:173-180; ghost cells are derived on the device)
__global__ void __launch_bounds__(kT) k_scatter_phi(const __grid_constant__ LevelDev L, const int* __restrict__ origin,
                                                    const double* __restrict__ grid, int n1, int n2) {
  const int b = blockIdx.x;
  double* v = L.var[0] + (size_t)b * L.bs;
  const int ox = origin[3 * b] + 1, oy = origin[3 * b + 1] + 1, oz = origin[3 * b + 2] + 1;
  for (int idx = threadIdx.x; idx < L.bs; idx += blockDim.x) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    v[idx] = grid[((size_t)(oz + k) * n2 + (oy + j)) * n1 + (ox + i)];
  }
}

struct Nx { int n[3]; };
// interp_bc (:237-270) of the planes bc_x0 .. bc_z1 (:159-167), evaluated straight from the potential grid
__global__ void __launch_bounds__(kT) k_bc_planes(const __grid_constant__ LevelDev L, const Face* __restrict__ faces,
                                                  const double* __restrict__ grid, const __grid_constant__ Nx nx, const __grid_constant__ Geom G) {
  const Face F = faces[blockIdx.x];
  const int nc = L.nc, d = F.nb >> 1, ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
  const int l1 = (F.nb & 1) ? nx.n[d] - 2 : 0, l2 = l1 + 1;          // the two layers the plane averages (0-based)
  double* out = const_cast<double*>(L.bcpl) + (size_t)F.plane * nc * nc;
  for (int idx = threadIdx.x; idx < nc * nc; idx += blockDim.x) {
    const int t1 = idx % nc, t2 = idx / nc;
    const double xa = F.ra + ((t1 + 1) - 0.5) * L.dr[ta], xb = F.rb + ((t2 + 1) - 0.5) * L.dr[tb];
    const double fa = (xa - G.r_min[ta]) * G.inv_dr[ta], fb = (xb - G.r_min[tb]) * G.inv_dr[tb];
    const int ia = (int)ceil(fa), ib = (int)ceil(fb);
    const double la = ia - fa, lb = ib - fb;
    const double w[2][2] = {{la * lb, la * (1 - lb)}, {(1 - la) * lb, (1 - la) * (1 - lb)}};   // w[a][b]
    double s = 0.0;
#pragma unroll
    for (int bq = 0; bq < 2; ++bq)
#pragma unroll
      for (int aq = 0; aq < 2; ++aq) {
        int c[3];
        c[ta] = ia + aq - 1; c[tb] = ib + bq - 1;                     // 1-based plane index -> 0-based
        c[d] = l1; const double v1 = grid[((size_t)c[2] * nx.n[1] + c[1]) * nx.n[0] + c[0]];
        c[d] = l2; const double v2 = grid[((size_t)c[2] * nx.n[1] + c[1]) * nx.n[0] + c[0]];
        s = s + w[aq][bq] * (0.5 * (v1 + v2));
      }
    out[idx] = s;
  }
}

}  // namespace

struct Plan {
  int nx[3], M[3];
  cufftHandle fwd = 0, inv = 0;
  bool have_fwd = false, have_inv = false, have_kernel = false;
  double *d_grid = nullptr, *d_pad = nullptr;
  cufftDoubleComplex *d_spec = nullptr, *d_kspec = nullptr;
  size_t n_spec = 0;
};

Plan* create(const int nx[3], std::string& err) {
  const CufftApi& A = cufft_api();
  if (!A.err.empty()) { err = A.err; return nullptr; }
  Plan* p = new Plan();
  for (int d = 0; d < 3; ++d) { p->nx[d] = nx[d]; p->M[d] = 2 * nx[d]; }
  const size_t N = (size_t)nx[0] * nx[1] * nx[2], NP = (size_t)p->M[0] * p->M[1] * p->M[2];
  p->n_spec = (size_t)(p->M[0] / 2 + 1) * p->M[1] * p->M[2];
  bool ok = cudaMalloc(&p->d_grid, N * sizeof(double)) == cudaSuccess && cudaMalloc(&p->d_pad, NP * sizeof(double)) == cudaSuccess &&
            cudaMalloc(&p->d_spec, p->n_spec * sizeof(cufftDoubleComplex)) == cudaSuccess &&
            cudaMalloc(&p->d_kspec, p->n_spec * sizeof(cufftDoubleComplex)) == cudaSuccess;
  if (!ok) { err = "free-space solve: out of device memory"; destroy(p); return nullptr; }
  if (A.Plan3d(&p->fwd, p->M[2], p->M[1], p->M[0], CUFFT_D2Z) != CUFFT_SUCCESS) { err = "cufftPlan3d (D2Z) failed"; destroy(p); return nullptr; }
  p->have_fwd = true;
  if (A.Plan3d(&p->inv, p->M[2], p->M[1], p->M[0], CUFFT_Z2D) != CUFFT_SUCCESS) { err = "cufftPlan3d (Z2D) failed"; destroy(p); return nullptr; }
  p->have_inv = true;
  return p;
}
void destroy(Plan* p) {
  if (!p) return;
  const CufftApi& A = cufft_api();
  if (p->have_fwd) A.Destroy(p->fwd);
  if (p->have_inv) A.Destroy(p->inv);
  cudaFree(p->d_grid); cudaFree(p->d_pad); cudaFree(p->d_spec); cudaFree(p->d_kspec);
  delete p;
}
double* grid(Plan* p) { return p->d_grid; }
size_t grid_size(const Plan* p) { return (size_t)p->nx[0] * p->nx[1] * p->nx[2]; }

bool set_kernel(Plan* p, int n_gauss, int n_range, const double* w, const double* kx, const double* ky, const double* kz,
                cudaStream_t s, std::string& err) {
  const CufftApi& A = cufft_api();
  const int nr1 = n_range + 1;
  const size_t nk = (size_t)n_gauss * nr1;
  double* d = nullptr;
  if (cudaMalloc(&d, (n_gauss + 3 * nk) * sizeof(double)) != cudaSuccess) { err = "free-space kernel: out of device memory"; return false; }
  double *dw = d, *dkx = d + n_gauss, *dky = dkx + nk, *dkz = dky + nk;
  bool ok = cudaMemcpyAsync(dw, w, n_gauss * sizeof(double), cudaMemcpyHostToDevice, s) == cudaSuccess &&
            cudaMemcpyAsync(dkx, kx, nk * sizeof(double), cudaMemcpyHostToDevice, s) == cudaSuccess &&
            cudaMemcpyAsync(dky, ky, nk * sizeof(double), cudaMemcpyHostToDevice, s) == cudaSuccess &&
            cudaMemcpyAsync(dkz, kz, nk * sizeof(double), cudaMemcpyHostToDevice, s) == cudaSuccess;
  if (ok) {
    const size_t NP = (size_t)p->M[0] * p->M[1] * p->M[2];
    k_build_kernel<<<grid_for(NP), kT, 0, s>>>(p->d_pad, p->M[0], p->M[1], p->M[2], p->nx[0], p->nx[1], p->nx[2], n_gauss, nr1, dw, dkx, dky, dkz);
    ok = A.SetStream(p->fwd, s) == CUFFT_SUCCESS && A.ExecD2Z(p->fwd, p->d_pad, p->d_kspec) == CUFFT_SUCCESS;
  }
  ok = ok && cudaStreamSynchronize(s) == cudaSuccess;     // host arrays / the staging allocation must outlive the work
  cudaFree(d);
  if (!ok) { err = "free-space kernel: transform failed"; return false; }
  p->have_kernel = true;
  return true;
}

bool convolve(Plan* p, double scal, cudaStream_t s, std::string& err) {
  const CufftApi& A = cufft_api();
  if (!p->have_kernel) { err = "free-space solve: no kernel"; return false; }
  const size_t N = grid_size(p), NP = (size_t)p->M[0] * p->M[1] * p->M[2];
  k_pad<<<grid_for(NP), kT, 0, s>>>(p->d_pad, p->d_grid, p->M[0], p->M[1], p->M[2], p->nx[0], p->nx[1], p->nx[2]);
  if (A.SetStream(p->fwd, s) != CUFFT_SUCCESS || A.ExecD2Z(p->fwd, p->d_pad, p->d_spec) != CUFFT_SUCCESS) { err = "cufftExecD2Z failed"; return false; }
  k_mul_spec<<<grid_for(p->n_spec), kT, 0, s>>>(p->d_spec, p->d_kspec, p->n_spec);
  if (A.SetStream(p->inv, s) != CUFFT_SUCCESS || A.ExecZ2D(p->inv, p->d_spec, p->d_pad) != CUFFT_SUCCESS) { err = "cufftExecZ2D failed"; return false; }
  k_crop<<<grid_for(N), kT, 0, s>>>(p->d_grid, p->d_pad, p->M[0], p->M[1], p->nx[0], p->nx[1], p->nx[2], scal / (double)NP);
  if (cudaGetLastError() != cudaSuccess) { err = "free-space solve: kernel launch failed"; return false; }
  return true;
}

void gather_rhs(const LevelDev& L, const int* origin, double rhs_fac, double* grid, const int nx[3], cudaStream_t s) {
  if (L.nb) k_gather_rhs<<<L.nb, kT, 0, s>>>(L, origin, rhs_fac, grid, nx[0], nx[1]);
}
void scatter_phi(const LevelDev& L, const int* origin, const double* grid, const int nx[3], cudaStream_t s) {
  if (L.nb) k_scatter_phi<<<L.nb, kT, 0, s>>>(L, origin, grid, nx[0], nx[1]);
}
void bc_planes(const LevelDev& L, const Face* faces, int n_faces, const double* grid, const int nx[3], const Geom& g, cudaStream_t s) {
  Nx n; n.n[0] = nx[0]; n.n[1] = nx[1]; n.n[2] = nx[2];
  if (n_faces) k_bc_planes<<<n_faces, kT, 0, s>>>(L, faces, grid, n, g);
}

}  // namespace omgfs
