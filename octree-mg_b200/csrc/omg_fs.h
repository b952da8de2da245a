// omg_fs.h -- device side of the free-space front end (mg_poisson_free_3d, src/m_free_space.f90:36-214): the coarse
// free-space solve as a zero-padded cuFFT convolution with the kernel of 1/r, and the kernels that move data between the
// box tree and the dense (n+2)^3 grid of the FFT level.  Plain pointers only; omg_core.cu holds the handle-side glue.
#pragma once
#include <string>

#include <cuda_runtime.h>

#include "omg_device.cuh"

namespace omgfs {

struct Plan;                                   // cuFFT plans, padded work arrays, kernel spectrum
Plan* create(const int nx[3], std::string& err);
void destroy(Plan* p);
// K(i, j, k) = sum_g w[g] kx[g][|i|] ky[g][|j|] kz[g][|k|] (n_range + 1 entries per Gaussian and direction), assembled on
// the padded grid with wrap-around and transformed once; Gaussians are added from the last to the first as the
// reference's Free_Kernel does (poisson_3d_fft/build_kernel.f90:1009-1050)
bool set_kernel(Plan* p, int n_gauss, int n_range, const double* w, const double* kx, const double* ky, const double* kz,
                cudaStream_t s, std::string& err);
double* grid(Plan* p);                         // dense (nx1, nx2, nx3) grid on the device, first index fastest
size_t grid_size(const Plan* p);
bool convolve(Plan* p, double scal, cudaStream_t s, std::string& err);      // grid <- scal * (K conv grid)

// boxes of the FFT level <-> dense grid (origin[3 b + d] = 0-based cell offset of box b in the grid without its ghost layer)
void gather_rhs(const LevelDev& L, const int* origin, double rhs_fac, double* grid, const int nx[3], cudaStream_t s);
void scatter_phi(const LevelDev& L, const int* origin, const double* grid, const int nx[3], cudaStream_t s);

// one physical-boundary face of a box: bc plane `plane` of the level, face nb (0-based), tangential coordinates of the
// box's low corner
struct Face { int plane, nb; double ra, rb; };
struct Geom { double r_min[3], inv_dr[3]; };   // of the FFT grid: r_min - dr / 2 and 1 / dr per direction (:125-137)
// ghost_cells_free_bc + interp_bc (:217-270): Dirichlet plane of every listed face by bilinear interpolation in the
// boundary planes 0.5 * (ghost layer + first layer) of the potential
void bc_planes(const LevelDev& L, const Face* faces, int n_faces, const double* grid, const int nx[3], const Geom& g, cudaStream_t s);

}  // namespace omgfs
