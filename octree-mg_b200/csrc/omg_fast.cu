// omg_fast.cu -- tuned sm_100a kernels for levels of box size 16 or 32 (any face kind: FK_SAME / FK_PHYS ghost
// layers are staged or evaluated in shared memory, FK_PLANE faces -- refinement boundaries, faces of another rank --
// come from the level's plane pool).  One CTA per tile = (box, slab of KT k-planes).  The tile and its ghost
// layer are staged in shared memory with TMA bulk copies (cp.async.bulk, completion on an mbarrier):
// thanks to the colour-split layout a colour half-sweep needs exactly the OTHER colour of the box, which
// is one contiguous 16 KiB block, plus 1 KiB (z) / 16 x 64 B (y) face slices of the neighbours' blocks.
// The arithmetic is the same expression, in the same order, as the general path (omg_ops.cuh), so both
// paths agree bitwise with each other and with the oracle.
#include "omg_fast.h"

#include <cstdint>
#include <cstdlib>
#include <type_traits>

namespace {

constexpr int kThreads = 256;
#ifndef OMG_VC_MINB
#define OMG_VC_MINB 3      // resident CTAs per SM the one-coefficient smoother is compiled for
#endif

#ifndef OMG_HOST_EMU   // tests/device_host/fast_emu.cpp supplies host stand-ins for this block and the three macros below
// ---- mbarrier / TMA bulk-copy PTX ---------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// raise the expected transaction bytes without arriving (the phase stays open until the final arrive)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
// global -> shared bulk copy (TMA, SASS UBLKCP); 16-byte aligned, size a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// shared -> global bulk copy (TMA store), bulk-group completion: the issuing thread commits the group and waits until the
// source has been read (the CTA may then exit; the writes themselves complete before the grid does)
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// wait for a peer's flag (same protocol as omg_comm.cu): an acquire load, so that the peer's puts -- released by its
// st.release.sys of the flag -- are visible afterwards; bounded by timeout_ns (0 = wait for ever): a peer that does not
// arrive raises the error word, which the host finds at its next synchronisation point (the handle is then poisoned)
__device__ __forceinline__ bool wait_flag(const unsigned long long* f, unsigned long long want, int* err, unsigned long long timeout_ns, int code) {
  unsigned long long t0 = 0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
    if (v >= want) return true;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (timeout_ns && t1 - t0 > timeout_ns) { *(volatile int*)err = code; __threadfence_system(); return false; }
    __nanosleep(64);
  }
}
__device__ __forceinline__ void signal_flag(unsigned long long* f, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_plane(const double* p) { return __ldcg(p); }   // peer-written planes: L2, never the read-only path
// ---- pieces of the persistent ring kernel (k_smooth_ring) ----
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {      // release.cta: the arriving thread's earlier writes are visible after the wait
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 8-byte asynchronous global -> shared copy (LDGSTS); completion is reported to an mbarrier by cp_async_arrive
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// generic-proxy writes to shared memory (ghost fix-ups) before the stage is handed back to the async proxy (TMA)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// one arrival on the mbarrier once all cp.async of this thread issued so far have landed (noinc: counted in the init count)
__device__ __forceinline__ void cp_async_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// barrier among the consumer warps only (barrier 1; the producer warp never joins)
__device__ __forceinline__ void consumer_sync(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory"); }
__device__ __forceinline__ int shfl_int(int v, int src) { return __shfl_sync(0xffffffffu, v, src); }
#define OMG_DYN_SMEM(name, al) extern __shared__ __align__(al) unsigned char name[]
#define OMG_LAUNCH(kern, grid, smem, stream, ...) kern<<<(grid), kThreads, (smem), (stream)>>>(__VA_ARGS__)
#define OMG_LAUNCH_T(kern, grid, threads, smem, stream, ...) kern<<<(grid), (threads), (smem), (stream)>>>(__VA_ARGS__)
#define OMG_PDL_ASM() do { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); asm volatile("griddepcontrol.wait;" ::: "memory"); } while (0)
#endif

// Programmatic dependent launch (option "pdl", off by default): the next kernel of the stream may become resident while
// this grid's last wave drains; it then blocks in pdl_sync() until this grid has completed and its stores are visible.
// Everything before pdl_sync() may only touch data no kernel writes (index / face tables, parameters, barrier set-up).
__device__ __forceinline__ void pdl_sync(int pdl) {
  if (pdl) OMG_PDL_ASM();
}

__device__ __forceinline__ void bc_coeffs(int type, double dr, int f, double& c0, double& c1, double& c2) {
  // bc_to_gc, src/m_ghost_cells.f90:682-698
  if (type == -10) { c0 = 2; c1 = -1; c2 = 0; }
  else if (type == -11) { c0 = dr * ((f & 1) ? 1.0 : -1.0); c1 = 1; c2 = 0; }
  else { c0 = 0; c1 = 2; c2 = -1; }
}

// In-kernel halo exchange, sending side (PutDev, omg_device.cuh): a CTA that has just written box `boxoff` (slab k0 .. k0 + KT of
// it) stores the box's face layers into the neighbour ranks' plane pools -- the copy they are not reading, (epoch + 1) & 1 --
// over NVLink; the last such CTA of the launch (n_bnd * tiles of them) signals the peers and bumps the epoch.  Used by the
// GSRB half-sweep kernels and by the prolongation kernel (whose put replaces the stand-alone ghost exchange that would
// follow it).  Called by every thread of the CTA.
template <int NC, int KT>
__device__ __forceinline__ void put_faces(const LevelDev& L, int slot, size_t boxoff, int k0, int tiles, int tid, int mode) {
  constexpr int HX = NC / 2;
  __syncthreads();
  const unsigned long long e = *L.put.epoch;
  const double* Abox = L.var[0] + boxoff;
  for (int it = L.put.ptr[slot]; it < L.put.ptr[slot + 1]; ++it) {
    const int f = L.put.item[3 * it], pr = L.put.item[3 * it + 1], pl = L.put.item[3 * it + 2];
    double* o = L.put.peer_base[pr] + (size_t)((e + 1ull) & 1ull) * (size_t)L.put.peer_stride[pr] + (size_t)pl * NC * NC;
    const int layer = (f & 1) ? 0 : NC - 1;          // f = the RECEIVER's face (buffer_for_nb, src/m_ghost_cells.f90:348-383)
    if (f < 4) {
      for (int r = tid; r < KT * NC; r += kThreads) {
        const int k = k0 + r / NC, t1 = r % NC;
        const int i = f < 2 ? layer : t1, j = f < 2 ? t1 : layer;
        o[k * NC + t1] = __ldcg(Abox + ((i + j + k) & 1) * L.cs + (k * NC + j) * HX + (i >> 1));
      }
    } else if (layer >= k0 && layer < k0 + KT) {
      for (int r = tid; r < NC * NC; r += kThreads) {
        const int i = r % NC, j = r / NC;
        o[r] = __ldcg(Abox + ((i + j + layer) & 1) * L.cs + (layer * NC + j) * HX + (i >> 1));
      }
    }
  }
  // mode 1: every thread fences its own remote stores before the CTA barrier; mode 2 (option "put_fence" = 2): the barrier
  // orders the CTA's stores before thread 0, whose system-scope fence + atomic is the (cumulative) release -- one fence per CTA
  if (mode != 2) __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    if (mode == 2) __threadfence_system();
    const unsigned int prev = atomicAdd(L.put.done, 1u);
    if (prev == (unsigned int)(L.put.n_bnd * tiles) - 1u) {
      *L.put.done = 0u;
      __threadfence_system();
      if (*(volatile int*)L.put.err == 0) {      // after a failed wait the exchange stops: no signal, epoch left alone
        for (int q = 0; q < L.put.n_peers; ++q)
          signal_flag(L.put.peer_flags[L.put.peers[q]] + L.put.channel * L.put.nranks + L.put.me, e + 1ull);
        *L.put.epoch = e + 1ull;
      }
    }
  }
}

// =================================================================================================
// GSRB colour half-sweep, Laplacian / Helmholtz (box_gs_lpl src/m_laplacian.f90:52-114, box_gs_helmh
// src/m_helmholtz.f90:49-108) over all boxes of a level.
//   reads : other colour of the tile (+ ghost layer), rhs of this colour      writes: this colour
//   algorithmic bytes: 12 B per cell of the level per launch
// Shared memory per CTA (NC=16, KT=16): 16 KiB tile + 6 KiB ghost layer.
// =================================================================================================
//
// VC = true: the variable-coefficient operators (box_gs_vlpl src/m_vlaplacian.f90:51-128, box_gs_vhelmh
// src/m_vhelmholtz.f90:61-141): the other colour of eps is staged the same way (its ghost layer comes from the
// stored-face pool: eps ghost cells are host data), eps of the updated cells is prefetched like rhs, and the six
// harmonic-mean face coefficients are recomputed per cell (6 fp64 divisions; precomputed face arrays would cost
// 24 B per cell and half-sweep instead of 8).  Algorithmic bytes: 12 + 8 (eps, both colours) per cell and launch.
template <int NC, int KT>
struct TileSmem {
  static constexpr int HX = NC / 2;
  double B[KT][NC][HX];     // other colour, dense as in global memory
  double zlo[NC][HX], zhi[NC][HX];
  double ylo[KT][HX], yhi[KT][HX];
  double xg[KT][NC];        // the single x ghost each row needs (low side if the row's updated cells start at i=0, else high side)
};
struct NoSmem {};
// NE = number of coefficient arrays: 0 (Laplacian / Helmholtz), 1 (vlaplacian / vhelmholtz: eps = variable 5 for all
// three directions), 3 (ahelmholtz, box_gs_ahelmh src/m_ahelmholtz.f90:69-162: variables 5, 6, 7 = eps of the x, y, z
// faces).  Array e is staged like phi; with NE = 3 array d only ever needs its ghost cells in direction d.
template <int NC, int KT, int NE>
struct SmoothSmem : TileSmem<NC, KT> {
  typename std::conditional<(NE > 0), TileSmem<NC, KT>[(NE > 0 ? NE : 1)], NoSmem>::type E;   // the same tile of eps
  uint64_t bar;
};

template <int NC, int KT, int NE>
__global__ void __launch_bounds__(kThreads, NE == 3 ? (KT == NC && NC == 16 ? 2 : 4) : NE ? (KT < NC && NC == 16 ? 5 : OMG_VC_MINB) : (NC == 16 ? 5 : 4)) k_smooth_lpl_fast(const __grid_constant__ LevelDev L,
                                                                          const __grid_constant__ ParamsDev P, int colour,
                                                                          const int* __restrict__ boxlist, int fused_put, int pdl) {
  constexpr int HX = NC / 2;
  constexpr int TILES = NC / KT;
  OMG_DYN_SMEM(smem_raw, 128);
  constexpr bool VC = NE > 0;
  constexpr int EY = NE == 3 ? 1 : 0, EZ = NE == 3 ? 2 : 0;      // which array serves the y / z faces (x: array 0)
  SmoothSmem<NC, KT, NE>& S = *reinterpret_cast<SmoothSmem<NC, KT, NE>*>(smem_raw);
  const int tid = threadIdx.x;

  const int b = boxlist ? boxlist[blockIdx.x / TILES] : blockIdx.x / TILES;   // a subset of the level (partition-face / interior boxes)
  // in-kernel halo exchange (PutDev, omg_device.cuh): is this box on a partition face?
  const int put_slot = fused_put ? L.put.slot_of[b] : -1;
  const bool bnd_put = put_slot >= 0;
  const int k0 = (blockIdx.x % TILES) * KT;
  const int oc = colour ^ 1;
  const double* __restrict__ phi = L.var[0];
  const size_t boxoff = (size_t)b * L.bs;
  const double* __restrict__ Bg = phi + boxoff + (size_t)oc * L.cs;     // other colour of this box
  double* __restrict__ Ag = L.var[0] + boxoff + (size_t)colour * L.cs;  // colour being updated
  const double* __restrict__ Rg = L.var[1] + boxoff + (size_t)colour * L.cs;

  // the tile itself depends on nothing but the box index: its copy starts before the face table is read
  if (tid == 0) mbar_init(&S.bar, 1);
  pdl_sync(pdl);
  if (tid == 0) {
    mbar_expect_tx(&S.bar, (1 + NE) * KT * NC * HX * 8);
    bulk_g2s(&S.B[0][0][0], Bg + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
    if constexpr (VC) {
#pragma unroll
      for (int e = 0; e < NE; ++e) bulk_g2s(&S.E[e].B[0][0][0], L.var[4 + e] + boxoff + (size_t)oc * L.cs + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
    }
  }
  // rhs (and eps) of this thread's cells (in flight while the bulk copies land)
  constexpr int PER = KT * NC * HX / kThreads;
  // COLUMN = true (box 16, constant coefficients): a thread owns one (m, j) column and KP consecutive k-planes, as in
  // k_smooth_ring (neighbour streams resolved once per thread, the centre value reused along k: 4 loads per cell)
  constexpr bool COLUMN = !VC && (kThreads % (NC * HX) == 0) && (KT % (kThreads / (NC * HX)) == 0) && ((KT / (kThreads / (NC * HX))) % 2 == 0);
  constexpr int COLS = NC * HX, KP = COLUMN ? KT / (kThreads / COLS) : 1;
  const int col = tid % COLS, kb = COLUMN ? (tid / COLS) * KP : 0;
  double rhs[PER], eps0[VC ? NE : 1][VC ? PER : 1];
  if constexpr (COLUMN) {
    static_assert(!COLUMN || KP == PER, "one cell per plane and thread");
#pragma unroll
    for (int q = 0; q < PER; ++q) rhs[q] = Rg[(size_t)(k0 + kb + q) * COLS + col];
  } else {
#pragma unroll
    for (int q = 0; q < PER; ++q) rhs[q] = Rg[(size_t)k0 * NC * HX + q * kThreads + tid];
  }
  if constexpr (VC) {
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      const double* __restrict__ Eg0 = L.var[4 + e] + boxoff + (size_t)colour * L.cs;
#pragma unroll
      for (int q = 0; q < PER; ++q) eps0[e][q] = Eg0[(size_t)k0 * NC * HX + q * kThreads + tid];
    }
  }

  int kind[6], arg[6];
#pragma unroll
  for (int f = 0; f < 6; ++f) { kind[f] = L.fkind[b * 6 + f]; arg[f] = L.farg[b * 6 + f]; }
  // z faces of an interior slab come from the box itself
  const bool zlo_in = (k0 > 0), zhi_in = (k0 + KT < NC);
  // the peers' faces of the previous half-sweep must have landed in the plane pool before they are read below
  if (bnd_put && tid >= 64 && tid < 64 + L.put.n_peers)
    wait_flag(L.put.flags + L.put.channel * L.put.nranks + L.put.peers[tid - 64], *L.put.epoch, L.put.err, L.put.timeout_ns,
              1 + L.put.peers[tid - 64] + 64 * L.put.channel);
  __syncthreads();

  if (tid < 32) {
    // ---- warp 0: the TMA bulk copies of the ghost layer ----
    if (tid == 0) {
      uint32_t bytes = 0;
      if (zlo_in || kind[4] == FK_SAME) bytes += NC * HX * 8;
      if (zhi_in || kind[5] == FK_SAME) bytes += NC * HX * 8;
      if (kind[2] == FK_SAME) bytes += KT * HX * 8;
      if (kind[3] == FK_SAME) bytes += KT * HX * 8;
      if (VC && zlo_in) bytes += NC * HX * 8;
      if (VC && zhi_in) bytes += NC * HX * 8;
      mbar_arrive_expect_tx(&S.bar, bytes);
    }
    __syncwarp();
    if constexpr (VC) {
      const double* __restrict__ Eg = L.var[4 + EZ] + boxoff + (size_t)oc * L.cs;
      if (tid == 3 && zlo_in) bulk_g2s(&S.E[EZ].zlo[0][0], Eg + (size_t)(k0 - 1) * NC * HX, NC * HX * 8, &S.bar);
      if (tid == 4 && zhi_in) bulk_g2s(&S.E[EZ].zhi[0][0], Eg + (size_t)(k0 + KT) * NC * HX, NC * HX * 8, &S.bar);
    }
    if (tid == 1) {
      if (zlo_in) bulk_g2s(&S.zlo[0][0], Bg + (size_t)(k0 - 1) * NC * HX, NC * HX * 8, &S.bar);
      else if (kind[4] == FK_SAME) bulk_g2s(&S.zlo[0][0], phi + (size_t)arg[4] * L.bs + (size_t)oc * L.cs + (size_t)(NC - 1) * NC * HX, NC * HX * 8, &S.bar);
    }
    if (tid == 2) {
      if (zhi_in) bulk_g2s(&S.zhi[0][0], Bg + (size_t)(k0 + KT) * NC * HX, NC * HX * 8, &S.bar);
      else if (kind[5] == FK_SAME) bulk_g2s(&S.zhi[0][0], phi + (size_t)arg[5] * L.bs + (size_t)oc * L.cs, NC * HX * 8, &S.bar);
    }
    // y faces: one 8*HX-byte row per k-plane from the neighbour's row j = NC-1 (low) / j = 0 (high)
    for (int q = tid; q < 2 * KT; q += 32) {
      const int hi = q / KT, kk = q % KT;
      if (kind[2 + hi] == FK_SAME) {
        const double* src = phi + (size_t)arg[2 + hi] * L.bs + (size_t)oc * L.cs + ((size_t)(k0 + kk) * NC + (hi ? 0 : NC - 1)) * HX;
        bulk_g2s(hi ? &S.yhi[kk][0] : &S.ylo[kk][0], src, HX * 8, &S.bar);
      }
    }
  } else {
    // ---- other warps: the x ghosts (one 8-byte value per row, strided in the neighbour) ----
    for (int r = tid - 32; r < KT * NC; r += kThreads - 32) {
      const int kk = r / NC, j = r % NC;
      const int sA = colour ^ ((j + k0 + kk) & 1);     // 0: this row's updated cells are i = 0, 2, ..: needs i = -1
      const int f = sA ? 1 : 0;
      if (kind[f] == FK_SAME)
        S.xg[kk][j] = phi[(size_t)arg[f] * L.bs + (size_t)oc * L.cs + ((size_t)(k0 + kk) * NC + j) * HX + (sA ? 0 : HX - 1)];
    }
  }

  // ---- ghost planes materialised elsewhere (refinement boundaries, faces owned by another rank): plain loads
  // of the other-colour entries of the nc x nc plane ----
  bool any_plane = false;
#pragma unroll
  for (int f = 0; f < 6; ++f) any_plane = any_plane || (kind[f] == FK_PLANE);
  if (any_plane) {
    const double* gpl = gpl_cur(L);
    if (!zlo_in && kind[4] == FK_PLANE)
      for (int r = tid; r < NC * HX; r += kThreads) { const int j = r / HX, m = r % HX; S.zlo[j][m] = ld_plane(gpl + (size_t)arg[4] * NC * NC + j * NC + 2 * m + ((oc ^ (j + 1)) & 1)); }
    if (!zhi_in && kind[5] == FK_PLANE)
      for (int r = tid; r < NC * HX; r += kThreads) { const int j = r / HX, m = r % HX; S.zhi[j][m] = ld_plane(gpl + (size_t)arg[5] * NC * NC + j * NC + 2 * m + ((oc ^ j) & 1)); }
    if (kind[2] == FK_PLANE)
      for (int r = tid; r < KT * HX; r += kThreads) { const int kk = r / HX, m = r % HX, k = k0 + kk; S.ylo[kk][m] = ld_plane(gpl + (size_t)arg[2] * NC * NC + k * NC + 2 * m + ((oc ^ (k + 1)) & 1)); }
    if (kind[3] == FK_PLANE)
      for (int r = tid; r < KT * HX; r += kThreads) { const int kk = r / HX, m = r % HX, k = k0 + kk; S.yhi[kk][m] = ld_plane(gpl + (size_t)arg[3] * NC * NC + k * NC + 2 * m + ((oc ^ k) & 1)); }
    if (kind[0] == FK_PLANE || kind[1] == FK_PLANE)
      for (int r = tid; r < KT * NC; r += kThreads) {
        const int kk = r / NC, j = r % NC, k = k0 + kk;
        const int f = (colour ^ ((j + k) & 1)) ? 1 : 0;
        if (kind[f] == FK_PLANE) S.xg[kk][j] = ld_plane(gpl + (size_t)arg[f] * NC * NC + k * NC + j);
      }
  }

  // ---- ghost layer of eps: the other-colour entries of the stored ghost faces (host data, any face kind) ----
  if constexpr (VC) {
    const double* __restrict__ fc = L.faces[4] + (size_t)b * 6 * NC * NC;
    const double* __restrict__ fcy = L.faces[4 + EY] + (size_t)b * 6 * NC * NC;
    const double* __restrict__ fcz = L.faces[4 + EZ] + (size_t)b * 6 * NC * NC;
    if (!zlo_in)
      for (int r = tid; r < NC * HX; r += kThreads) { const int j = r / HX, m = r % HX; S.E[EZ].zlo[j][m] = fcz[4 * NC * NC + j * NC + 2 * m + ((oc ^ (j + 1)) & 1)]; }
    if (!zhi_in)
      for (int r = tid; r < NC * HX; r += kThreads) { const int j = r / HX, m = r % HX; S.E[EZ].zhi[j][m] = fcz[5 * NC * NC + j * NC + 2 * m + ((oc ^ j) & 1)]; }
    for (int r = tid; r < KT * HX; r += kThreads) {
      const int kk = r / HX, m = r % HX, k = k0 + kk;
      S.E[EY].ylo[kk][m] = fcy[2 * NC * NC + k * NC + 2 * m + ((oc ^ (k + 1)) & 1)];
      S.E[EY].yhi[kk][m] = fcy[3 * NC * NC + k * NC + 2 * m + ((oc ^ k) & 1)];
    }
    for (int r = tid; r < KT * NC; r += kThreads) {
      const int kk = r / NC, j = r % NC, k = k0 + kk;
      const int f = (colour ^ ((j + k) & 1)) ? 1 : 0;
      S.E[0].xg[kk][j] = fc[f * NC * NC + k * NC + j];
    }
  }

  mbar_wait(&S.bar, 0);
  __syncthreads();

  // ---- physical boundaries: ghost = c0*b + c1*phi(1) + c2*phi(2), evaluated like bc_to_gc ----
  bool any_phys = false;
#pragma unroll
  for (int f = 0; f < 6; ++f) any_phys = any_phys || (kind[f] == FK_PHYS);
  if (any_phys) {
    for (int f = 0; f < 6; ++f) {
      if (kind[f] != FK_PHYS) continue;
      if (f == 4 && zlo_in) continue;
      if (f == 5 && zhi_in) continue;
      int type = L.ftype[b * 6 + f];
      if (type == 0) type = P.bc_type[f];
      double c0, c1, c2;
      bc_coeffs(type, L.dr[f >> 1], f, c0, c1, c2);
      const double* pl = arg[f] >= 0 ? L.bcpl + (size_t)arg[f] * NC * NC : nullptr;
      const double bconst = P.bc_value[f];
      if (f < 2) {
        for (int r = tid; r < KT * NC; r += kThreads) {
          const int kk = r / NC, j = r % NC, k = k0 + kk;
          const int sA = colour ^ ((j + k) & 1);
          if (sA != f) continue;                       // this row needs the other side
          const int m = f ? HX - 1 : 0;
          const double bv = pl ? pl[k * NC + j] : bconst;
          S.xg[kk][j] = c0 * bv + c1 * Ag[((size_t)k * NC + j) * HX + m] + c2 * S.B[kk][j][m];
        }
      } else if (f < 4) {
        const int j1 = (f & 1) ? NC - 1 : 0, j2 = (f & 1) ? NC - 2 : 1;
        for (int r = tid; r < KT * HX; r += kThreads) {
          const int kk = r / HX, m = r % HX, k = k0 + kk;
          const int i = 2 * m + (colour ^ ((j1 + k) & 1));
          const double bv = pl ? pl[k * NC + i] : bconst;
          const double g = c0 * bv + c1 * Ag[((size_t)k * NC + j1) * HX + m] + c2 * S.B[kk][j2][m];
          if (f & 1) S.yhi[kk][m] = g; else S.ylo[kk][m] = g;
        }
      } else {
        const int k1 = (f & 1) ? NC - 1 : 0, kk1 = k1 - k0, kk2 = (f & 1) ? kk1 - 1 : kk1 + 1;
        for (int r = tid; r < NC * HX; r += kThreads) {
          const int j = r / HX, m = r % HX;
          const int i = 2 * m + (colour ^ ((j + k1) & 1));
          const double bv = pl ? pl[j * NC + i] : bconst;
          const double g = c0 * bv + c1 * Ag[((size_t)k1 * NC + j) * HX + m] + c2 * S.B[kk2][j][m];
          if (f & 1) S.zhi[j][m] = g; else S.zlo[j][m] = g;
        }
      }
    }
    __syncthreads();
  }

  // ---- update this colour ----
  const double idx2 = L.idr2[0], idy2 = L.idr2[1], idz2 = L.idr2[2], fac = L.fac;
  if constexpr (COLUMN) {
    using TS = TileSmem<NC, KT>;
    constexpr int oB = offsetof(TS, B) / 8, oZl = offsetof(TS, zlo) / 8, oZh = offsetof(TS, zhi) / 8, oYl = offsetof(TS, ylo) / 8,
                  oYh = offsetof(TS, yhi) / 8, oXg = offsetof(TS, xg) / 8, PL = NC * HX;
    const int m = col % HX, j = col / HX;
    const int s0 = colour ^ ((j + k0 + kb) & 1);          // sA of the thread's first plane; alternates with kk
    const int x0 = m > 0 ? oB + kb * PL + j * HX + m - 1 : oXg + kb * NC + j, x0s = m > 0 ? PL : NC;
    const int x1 = m < HX - 1 ? oB + kb * PL + j * HX + m + 1 : oXg + kb * NC + j, x1s = m < HX - 1 ? PL : NC;
    const int xe = s0 ? x1 : x0, xes = s0 ? x1s : x0s, xo = s0 ? x0 : x1, xos = s0 ? x0s : x1s;
    const int yl = j > 0 ? oB + kb * PL + (j - 1) * HX + m : oYl + kb * HX + m, yls = j > 0 ? PL : HX;
    const int yh = j < NC - 1 ? oB + kb * PL + (j + 1) * HX + m : oYh + kb * HX + m, yhs = j < NC - 1 ? PL : HX;
    const int ctr = oB + kb * PL + j * HX + m;
    const int zl = kb > 0 ? ctr - PL : oZl + j * HX + m;
    const int zh = kb + KP < KT ? ctr + KP * PL : oZh + j * HX + m;
    const double* __restrict__ W = reinterpret_cast<const double*>(static_cast<const TS*>(&S));
    double* __restrict__ out = Ag + (size_t)(k0 + kb) * PL + col;
    double cm = W[zl], c0 = W[ctr];
#pragma unroll
    for (int q = 0; q < KP; ++q) {
      const double cp = q < KP - 1 ? W[ctr + (q + 1) * PL] : W[zh];
      const double xn = (q & 1) ? W[xo + q * xos] : W[xe + q * xes];
      const double ylo = W[yl + q * yls], yhi = W[yh + q * yhs];
      out[q * PL] = fac * (idx2 * (xn + c0) + idy2 * (yhi + ylo) + idz2 * (cp + cm) - rhs[q]);
      cm = c0; c0 = cp;
    }
  } else {
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int idx = q * kThreads + tid;
    const int m = idx % HX, j = (idx / HX) % NC, kk = idx / (HX * NC);
    const int sA = colour ^ ((j + k0 + kk) & 1);
    const int ml = m - 1 + sA, mh = m + sA;
    const double xlo = ml >= 0 ? S.B[kk][j][ml] : S.xg[kk][j];
    const double xhi = mh < HX ? S.B[kk][j][mh] : S.xg[kk][j];
    const double ylo = j > 0 ? S.B[kk][j - 1][m] : S.ylo[kk][m];
    const double yhi = j < NC - 1 ? S.B[kk][j + 1][m] : S.yhi[kk][m];
    const double zlo = kk > 0 ? S.B[kk - 1][j][m] : S.zlo[j][m];
    const double zhi = kk < KT - 1 ? S.B[kk + 1][j][m] : S.zhi[j][m];
    if constexpr (VC) {
      const double u[6] = {xlo, xhi, ylo, yhi, zlo, zhi};
      double a0[6], a[6];
#pragma unroll
      for (int f = 0; f < 6; ++f) a0[f] = eps0[NE == 3 ? (f >> 1) : 0][q];
      a[0] = ml >= 0 ? S.E[0].B[kk][j][ml] : S.E[0].xg[kk][j];
      a[1] = mh < HX ? S.E[0].B[kk][j][mh] : S.E[0].xg[kk][j];
      a[2] = j > 0 ? S.E[EY].B[kk][j - 1][m] : S.E[EY].ylo[kk][m];
      a[3] = j < NC - 1 ? S.E[EY].B[kk][j + 1][m] : S.E[EY].yhi[kk][m];
      a[4] = kk > 0 ? S.E[EZ].B[kk - 1][j][m] : S.E[EZ].zlo[j][m];
      a[5] = kk < KT - 1 ? S.E[EZ].B[kk + 1][j][m] : S.E[EZ].zhi[j][m];
      Ag[(size_t)k0 * NC * HX + idx] = P.coeff_math ? gs_vlpl_value_tol(L.idr2, u, a0, a, rhs[q], P.lambda, P.op != OP_VLPL)
                                                    : gs_vlpl_value(L.idr2, u, a0, a, rhs[q], P.lambda, P.op != OP_VLPL);
    } else {
      Ag[(size_t)k0 * NC * HX + idx] = fac * (idx2 * (xhi + xlo) + idy2 * (yhi + ylo) + idz2 * (zhi + zlo) - rhs[q]);
    }
  }
  }

  // ---- partition-face box: store the fresh face layers into the neighbour ranks' plane pools; the last such CTA of the
  // launch signals the peers and bumps the epoch (put_faces) ----
  if (bnd_put) put_faces<NC, KT>(L, put_slot, boxoff, k0, TILES, tid, fused_put);
}

// =================================================================================================
// Persistent ring version of the GSRB colour half-sweep (Laplacian / Helmholtz; same arithmetic, same bits as
// k_smooth_lpl_fast).  One CTA per SM walks the tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...  A producer warp
// keeps RING_STAGES tiles in flight: per tile one 16 KiB TMA bulk copy of the other colour, one of the rhs of the
// updated colour, bulk copies of the z / y ghost slices and 8-byte cp.async copies of the x ghosts, all completing on
// the stage's "full" mbarrier; eight consumer warps update tile n while the copies of tiles n+1 .. n+STAGES-1 are in
// flight and hand the stage back through its "empty" mbarrier.  This removes what bounded the one-tile-per-CTA kernel
// (load -> wait -> compute -> store in sequence, overlap only between co-resident CTAs, a partly empty last wave).
// Physical-boundary faces: the producer stages the box's OWN first layer (the phi(1) of bc_to_gc) in the ghost slot,
// the consumers turn it into the ghost value in place; plane faces (refinement boundaries, other ranks) are read from
// the plane pool by the consumers.  Both need one consumer-only barrier, paid by boundary tiles only.
// =================================================================================================
// NCONS consumer threads (+ 1 producer warp) per CTA, MINB CTAs per SM: the consumers' update is latency-bound per warp
// (profiles/r02c_ncu_smooth_ring.txt: 8 consumer warps on an SM issue 23 % of the time and the half-sweep takes 0.50 ms
// against 0.30 ms of the one-tile-per-CTA kernel), so an SM needs 16 .. 32 of them.
template <int NC, int KT>
struct RingStage {
  static constexpr int HX = NC / 2;
  double B[KT][NC][HX];     // other colour
  double R[KT][NC][HX];     // rhs of the updated colour
  double zlo[NC][HX], zhi[NC][HX];
  double ylo[KT][HX], yhi[KT][HX];
  double xg[KT][NC];
  int hdr[16];              // kind[6], arg[6], box, k0, put slot, spare: written by the producer before its arrive
};
template <int NC, int KT, int STAGES>
struct RingSmem {
  RingStage<NC, KT> st[STAGES];
  uint64_t full[STAGES], empty[STAGES];
};

template <int NC, int KT, int STAGES, int NCONS, int MINB, int NPW>
__global__ void __launch_bounds__(NCONS + 32 * NPW, MINB) k_smooth_ring(const __grid_constant__ LevelDev L, const __grid_constant__ ParamsDev P, int colour,
                                                              const int* __restrict__ boxlist, int n_tiles, int fused_put) {
  constexpr int HX = NC / 2;
  constexpr int TILES = NC / KT;
  constexpr int kRingConsumers = NCONS;
  constexpr int NPT = 32 * NPW;                  // producer threads: one warp issues 645 instructions per tile (ncu r02e) = 3900 cycles,
                                                 // more than the tile's share of the HBM time; NPW warps split the ghost gathers
  using SM = RingSmem<NC, KT, STAGES>;
  OMG_DYN_SMEM(smem_raw, 128);
  SM& S = *reinterpret_cast<SM*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int oc = colour ^ 1;
  const double* __restrict__ phi = L.var[0];
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&S.full[s], 1 + NPT); mbar_init(&S.empty[s], kRingConsumers / 32); }
  }
  __syncthreads();

  if (tid >= kRingConsumers) {
    // ------------------------------------------------ producer warp ------------------------------------------------
    // The face tables of a tile are read one tile ahead (box index of a box list: two ahead): a single warp that loaded
    // them only when it needed them paid one global-memory latency per tile, which bounded the whole kernel at one tile
    // per ~2.2 us and SM whatever the number of stages or consumers (profiles/r02d_*: 0.51 ms with one CTA per SM).
    int it = 0;
    const int stride = gridDim.x;
    const int pt = tid - kRingConsumers;           // 0 .. NPT-1; warp 0 of the producers issues the bulk copies
    auto box_of = [&](int tt) { return tt < n_tiles ? (boxlist ? boxlist[tt / TILES] : tt / TILES) : 0; };
    int b_n = box_of(blockIdx.x), b_nn = box_of(blockIdx.x + stride);
    int kv_n = 0, av_n = 0, slot_n = -1;
    auto fetch_tables = [&](int bb) {
      if (lane < 6) { kv_n = L.fkind[bb * 6 + lane]; av_n = L.farg[bb * 6 + lane]; }
      if (fused_put && lane == 0) slot_n = L.put.slot_of[bb];
    };
    fetch_tables(b_n);
    for (int t = blockIdx.x; t < n_tiles; t += stride, ++it) {
      const int s = it % STAGES, use = it / STAGES;
      const int b = b_n, kv = kv_n, av = av_n, slot_l = slot_n;
      b_n = b_nn; b_nn = box_of(t + 2 * stride);
      if (t + stride < n_tiles) fetch_tables(b_n);              // in flight while this tile's copies are issued
      if (use > 0) mbar_wait(&S.empty[s], (use - 1) & 1);       // the consumers have finished with the stage's previous tile
      RingStage<NC, KT>& T = S.st[s];
      const int k0 = (t % TILES) * KT;
      int kind[6], arg[6];
#pragma unroll
      for (int f = 0; f < 6; ++f) { kind[f] = shfl_int(kv, f); arg[f] = shfl_int(av, f); }
      int put_slot = -1;
      if (fused_put) put_slot = shfl_int(slot_l, 0);
      // partition-face box: the peers' faces of the previous half-sweep must have landed before the consumers read them
      if (put_slot >= 0 && pt < L.put.n_peers)
        wait_flag(L.put.flags + L.put.channel * L.put.nranks + L.put.peers[pt], *L.put.epoch, L.put.err, L.put.timeout_ns,
                  1 + L.put.peers[pt] + 64 * L.put.channel);
      __syncwarp();
      const size_t boxoff = (size_t)b * L.bs;
      const double* __restrict__ Bg = phi + boxoff + (size_t)oc * L.cs;        // other colour of this box
      const double* __restrict__ Ag = phi + boxoff + (size_t)colour * L.cs;    // colour being updated (its old values: phi(1) of bc_to_gc)
      const double* __restrict__ Rg = L.var[1] + boxoff + (size_t)colour * L.cs;
      const bool zlo_in = (k0 > 0), zhi_in = (k0 + KT < NC);
      if (pt == 0) {
#pragma unroll
        for (int f = 0; f < 6; ++f) { T.hdr[f] = kind[f]; T.hdr[6 + f] = arg[f]; }
        T.hdr[12] = b; T.hdr[13] = k0; T.hdr[14] = put_slot;
        uint32_t bytes = 2u * KT * NC * HX * 8;
        if (zlo_in || kind[4] != FK_PLANE) bytes += NC * HX * 8;
        if (zhi_in || kind[5] != FK_PLANE) bytes += NC * HX * 8;
        mbar_arrive_expect_tx(&S.full[s], bytes);
        bulk_g2s(&T.B[0][0][0], Bg + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.full[s]);
      }
      if (pt == 1) bulk_g2s(&T.R[0][0][0], Rg + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.full[s]);
      if (pt == 2) {
        if (zlo_in) bulk_g2s(&T.zlo[0][0], Bg + (size_t)(k0 - 1) * NC * HX, NC * HX * 8, &S.full[s]);
        else if (kind[4] == FK_SAME) bulk_g2s(&T.zlo[0][0], phi + (size_t)arg[4] * L.bs + (size_t)oc * L.cs + (size_t)(NC - 1) * NC * HX, NC * HX * 8, &S.full[s]);
        else if (kind[4] == FK_PHYS) bulk_g2s(&T.zlo[0][0], Ag, NC * HX * 8, &S.full[s]);                                  // own layer k = 0
      }
      if (pt == 3) {
        if (zhi_in) bulk_g2s(&T.zhi[0][0], Bg + (size_t)(k0 + KT) * NC * HX, NC * HX * 8, &S.full[s]);
        else if (kind[5] == FK_SAME) bulk_g2s(&T.zhi[0][0], phi + (size_t)arg[5] * L.bs + (size_t)oc * L.cs, NC * HX * 8, &S.full[s]);
        else if (kind[5] == FK_PHYS) bulk_g2s(&T.zhi[0][0], Ag + (size_t)(NC - 1) * NC * HX, NC * HX * 8, &S.full[s]);       // own layer k = NC-1
      }
      // y faces: one 8*HX-byte row per k-plane: the neighbour's row j = NC-1 (low) / j = 0 (high), or the box's own row
      // j = 0 / NC-1 at a physical boundary
      // (16-byte cp.async pieces: 32 + 32 short rows per tile would keep the TMA unit busier than the four big copies)
      constexpr int CH = HX / 2;                       // 16-byte pieces per row
      for (int q = pt; q < 2 * KT * CH; q += NPT) {
        const int c = q % CH, kk = (q / CH) % KT, hi = q / (CH * KT);
        const int kf = hi ? kind[3] : kind[2], af = hi ? arg[3] : arg[2];
        const double* src = nullptr;
        if (kf == FK_SAME) src = phi + (size_t)af * L.bs + (size_t)oc * L.cs + ((size_t)(k0 + kk) * NC + (hi ? 0 : NC - 1)) * HX;
        else if (kf == FK_PHYS) src = Ag + ((size_t)(k0 + kk) * NC + (hi ? NC - 1 : 0)) * HX;
        if (src) cp_async16((hi ? &T.yhi[kk][0] : &T.ylo[kk][0]) + 2 * c, src + 2 * c);
      }
      // x ghosts: one 8-byte value per row (strided in the neighbour / in the box itself)
      for (int r = pt; r < KT * NC; r += NPT) {
        const int kk = r / NC, j = r % NC;
        const int sA = colour ^ ((j + k0 + kk) & 1);     // 0: this row's updated cells are i = 0, 2, ..: needs i = -1
        const int kf = sA ? kind[1] : kind[0], af = sA ? arg[1] : arg[0];
        const size_t row = ((size_t)(k0 + kk) * NC + j) * HX;
        if (kf == FK_SAME) cp_async8(&T.xg[kk][j], phi + (size_t)af * L.bs + (size_t)oc * L.cs + row + (sA ? 0 : HX - 1));
        else if (kf == FK_PHYS) cp_async8(&T.xg[kk][j], Ag + row + (sA ? HX - 1 : 0));
      }
      cp_async_arrive(&S.full[s]);
    }
    return;
  }

  // -------------------------------------------------- consumer warps --------------------------------------------------
  const double idx2 = L.idr2[0], idy2 = L.idr2[1], idz2 = L.idr2[2], fac = L.fac;
  // Update loop: a thread owns one (m, j) column of the tile and KP consecutive k-planes of it.  Which neighbour lies
  // outside the tile then depends on the thread only, so the stage-relative offsets of the six neighbour streams are
  // computed once per kernel; and the other-colour value at (m, j) of plane kk is at once the z neighbour of planes
  // kk -+ 1 and one of the two x neighbours of plane kk, so it is loaded once and kept in registers: 5 shared-memory
  // loads per cell (rhs included) instead of 7.  (xhi + xlo and zhi + zlo are commutative: same bits.)
  constexpr int COLS = NC * HX, NG = kRingConsumers / COLS, KP = KT / NG;
  static_assert(kRingConsumers % COLS == 0 && KT % NG == 0 && KP % 2 == 0 && KT % 2 == 0, "consumer threads must tile the (m, j) columns");
  using ST = RingStage<NC, KT>;
  const int col = tid % COLS, m = col % HX, j = col / HX, kb = (tid / COLS) * KP;
  constexpr int oB = offsetof(ST, B) / 8, oR = offsetof(ST, R) / 8, oZl = offsetof(ST, zlo) / 8, oZh = offsetof(ST, zhi) / 8,
                oYl = offsetof(ST, ylo) / 8, oYh = offsetof(ST, yhi) / 8, oXg = offsetof(ST, xg) / 8, PL = NC * HX;
  const int s0 = colour ^ ((j + kb) & 1);               // sA of the thread's first plane (k0 is even); alternates with kk
  // x neighbour that is not the column itself: i - 1 where sA = 0, i + 1 where sA = 1
  const int x0 = m > 0 ? oB + kb * PL + j * HX + m - 1 : oXg + kb * NC + j, x0s = m > 0 ? PL : NC;
  const int x1 = m < HX - 1 ? oB + kb * PL + j * HX + m + 1 : oXg + kb * NC + j, x1s = m < HX - 1 ? PL : NC;
  const int xe = s0 ? x1 : x0, xes = s0 ? x1s : x0s, xo = s0 ? x0 : x1, xos = s0 ? x0s : x1s;   // even / odd planes of the thread
  const int yl = j > 0 ? oB + kb * PL + (j - 1) * HX + m : oYl + kb * HX + m, yls = j > 0 ? PL : HX;
  const int yh = j < NC - 1 ? oB + kb * PL + (j + 1) * HX + m : oYh + kb * HX + m, yhs = j < NC - 1 ? PL : HX;
  const int ctr = oB + kb * PL + j * HX + m;
  const int zl = kb > 0 ? ctr - PL : oZl + j * HX + m;
  const int zh = kb + KP < KT ? ctr + KP * PL : oZh + j * HX + m;
  const int rr = oR + kb * PL + j * HX + m;
  int it = 0;
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
    const int s = it % STAGES, use = it / STAGES;
    mbar_wait(&S.full[s], use & 1);
    RingStage<NC, KT>& T = S.st[s];
    int kind[6], arg[6];
#pragma unroll
    for (int f = 0; f < 6; ++f) { kind[f] = T.hdr[f]; arg[f] = T.hdr[6 + f]; }
    const int b = T.hdr[12], k0 = T.hdr[13], put_slot = T.hdr[14];
    const size_t boxoff = (size_t)b * L.bs;
    double* __restrict__ Ag = L.var[0] + boxoff + (size_t)colour * L.cs;
    const bool zlo_in = (k0 > 0), zhi_in = (k0 + KT < NC);
    bool fix = false;
#pragma unroll
    for (int f = 0; f < 6; ++f) fix = fix || (kind[f] != FK_SAME);
    if (fix) {
      // ---- ghost planes materialised elsewhere: the other-colour entries of the nc x nc plane ----
      const double* gpl = gpl_cur(L);
      if (!zlo_in && kind[4] == FK_PLANE)
        for (int r = tid; r < NC * HX; r += kRingConsumers) { const int j = r / HX, m = r % HX; T.zlo[j][m] = ld_plane(gpl + (size_t)arg[4] * NC * NC + j * NC + 2 * m + ((oc ^ (j + 1)) & 1)); }
      if (!zhi_in && kind[5] == FK_PLANE)
        for (int r = tid; r < NC * HX; r += kRingConsumers) { const int j = r / HX, m = r % HX; T.zhi[j][m] = ld_plane(gpl + (size_t)arg[5] * NC * NC + j * NC + 2 * m + ((oc ^ j) & 1)); }
      if (kind[2] == FK_PLANE)
        for (int r = tid; r < KT * HX; r += kRingConsumers) { const int kk = r / HX, m = r % HX, k = k0 + kk; T.ylo[kk][m] = ld_plane(gpl + (size_t)arg[2] * NC * NC + k * NC + 2 * m + ((oc ^ (k + 1)) & 1)); }
      if (kind[3] == FK_PLANE)
        for (int r = tid; r < KT * HX; r += kRingConsumers) { const int kk = r / HX, m = r % HX, k = k0 + kk; T.yhi[kk][m] = ld_plane(gpl + (size_t)arg[3] * NC * NC + k * NC + 2 * m + ((oc ^ k) & 1)); }
      if (kind[0] == FK_PLANE || kind[1] == FK_PLANE)
        for (int r = tid; r < KT * NC; r += kRingConsumers) {
          const int kk = r / NC, j = r % NC, k = k0 + kk;
          const int f = (colour ^ ((j + k) & 1)) ? 1 : 0;
          if ((f ? kind[1] : kind[0]) == FK_PLANE) T.xg[kk][j] = ld_plane(gpl + (size_t)(f ? arg[1] : arg[0]) * NC * NC + k * NC + j);
        }
      // ---- physical boundaries: ghost = c0*b + c1*phi(1) + c2*phi(2) (bc_to_gc); the slot holds phi(1) ----
#pragma unroll
      for (int f = 0; f < 6; ++f) {
        if (kind[f] != FK_PHYS) continue;
        if (f == 4 && zlo_in) continue;
        if (f == 5 && zhi_in) continue;
        int type = L.ftype[b * 6 + f];
        if (type == 0) type = P.bc_type[f];
        double c0, c1, c2;
        bc_coeffs(type, L.dr[f >> 1], f, c0, c1, c2);
        const double* pl = arg[f] >= 0 ? L.bcpl + (size_t)arg[f] * NC * NC : nullptr;
        const double bconst = P.bc_value[f];
        if (f < 2) {
          for (int r = tid; r < KT * NC; r += kRingConsumers) {
            const int kk = r / NC, j = r % NC, k = k0 + kk;
            const int sA = colour ^ ((j + k) & 1);
            if (sA != f) continue;                       // this row needs the other side
            const int m = f ? HX - 1 : 0;
            const double bv = pl ? pl[k * NC + j] : bconst;
            T.xg[kk][j] = c0 * bv + c1 * T.xg[kk][j] + c2 * T.B[kk][j][m];
          }
        } else if (f < 4) {
          const int j1 = (f & 1) ? NC - 1 : 0, j2 = (f & 1) ? NC - 2 : 1;
          for (int r = tid; r < KT * HX; r += kRingConsumers) {
            const int kk = r / HX, m = r % HX, k = k0 + kk;
            const int i = 2 * m + (colour ^ ((j1 + k) & 1));
            const double bv = pl ? pl[k * NC + i] : bconst;
            double& g = (f & 1) ? T.yhi[kk][m] : T.ylo[kk][m];
            g = c0 * bv + c1 * g + c2 * T.B[kk][j2][m];
          }
        } else {
          const int k1 = (f & 1) ? NC - 1 : 0, kk1 = k1 - k0, kk2 = (f & 1) ? kk1 - 1 : kk1 + 1;
          for (int r = tid; r < NC * HX; r += kRingConsumers) {
            const int j = r / HX, m = r % HX;
            const int i = 2 * m + (colour ^ ((j + k1) & 1));
            const double bv = pl ? pl[j * NC + i] : bconst;
            double& g = (f & 1) ? T.zhi[j][m] : T.zlo[j][m];
            g = c0 * bv + c1 * g + c2 * T.B[kk2][j][m];
          }
        }
      }
      consumer_sync(kRingConsumers);
    }

    // ---- update this colour ----
    {
      const double* __restrict__ W = reinterpret_cast<const double*>(&T);
      double* __restrict__ out = Ag + (size_t)(k0 + kb) * PL + col;
      double cm = W[zl], c0 = W[ctr];
#pragma unroll
      for (int q = 0; q < KP; ++q) {
        const double cp = q < KP - 1 ? W[ctr + (q + 1) * PL] : W[zh];
        const double xn = (q & 1) ? W[xo + q * xos] : W[xe + q * xes];
        const double ylo = W[yl + q * yls], yhi = W[yh + q * yhs];
        out[q * PL] = fac * (idx2 * (xn + c0) + idy2 * (yhi + ylo) + idz2 * (cp + cm) - W[rr + q * PL]);
        cm = c0; c0 = cp;
      }
    }

    // ---- partition-face box: store the fresh face layers into the neighbour ranks' plane pools (see k_smooth_lpl_fast) ----
    if (put_slot >= 0) {
      consumer_sync(kRingConsumers);
      const unsigned long long e = *L.put.epoch;
      const double* Abox = L.var[0] + boxoff;
      for (int q = L.put.ptr[put_slot]; q < L.put.ptr[put_slot + 1]; ++q) {
        const int f = L.put.item[3 * q], pr = L.put.item[3 * q + 1], pl = L.put.item[3 * q + 2];
        double* o = L.put.peer_base[pr] + (size_t)((e + 1ull) & 1ull) * (size_t)L.put.peer_stride[pr] + (size_t)pl * NC * NC;
        const int layer = (f & 1) ? 0 : NC - 1;          // f = the RECEIVER's face
        if (f < 4) {
          for (int r = tid; r < KT * NC; r += kRingConsumers) {
            const int k = k0 + r / NC, t1 = r % NC;
            const int i = f < 2 ? layer : t1, j = f < 2 ? t1 : layer;
            o[k * NC + t1] = __ldcg(Abox + ((i + j + k) & 1) * L.cs + (k * NC + j) * HX + (i >> 1));
          }
        } else if (layer >= k0 && layer < k0 + KT) {
          for (int r = tid; r < NC * NC; r += kRingConsumers) {
            const int i = r % NC, j = r / NC;
            o[r] = __ldcg(Abox + ((i + j + layer) & 1) * L.cs + (layer * NC + j) * HX + (i >> 1));
          }
        }
      }
      __threadfence_system();
      consumer_sync(kRingConsumers);
      if (tid == 0) {
        const unsigned int prev = atomicAdd(L.put.done, 1u);
        if (prev == (unsigned int)(L.put.n_bnd * TILES) - 1u) {
          *L.put.done = 0u;
          __threadfence_system();
          if (*(volatile int*)L.put.err == 0) {
            for (int q = 0; q < L.put.n_peers; ++q)
              signal_flag(L.put.peer_flags[L.put.peers[q]] + L.put.channel * L.put.nranks + L.put.me, e + 1ull);
            *L.put.epoch = e + 1ull;
          }
        }
      }
    }
    if (fix) fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(&S.empty[s]);
  }
}

// integer tuning knobs from the environment
inline int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e && *e ? atoi(e) : dflt;
}
// number of SMs of the current device (grid of the persistent kernels)
inline int sm_count() {
  static int n = 0;
  if (!n) {
#ifdef OMG_HOST_EMU
    n = 3;
#else
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
#endif
  }
  return n;
}
int g_ring = 1;          // option "ring": 1 = persistent ring smoother on levels with at least 4 tiles per SM, 2 = always, 0 = never

template <int NC, int KT, int STAGES, int NCONS, int MINB, int NPW>
bool launch_smooth_ring(const LevelDev& L, const ParamsDev& P, int colour, const int* list, int n, int fused, cudaStream_t s) {
  static bool configured = false;
  const size_t smem = sizeof(RingSmem<NC, KT, STAGES>);
  auto kern = k_smooth_ring<NC, KT, STAGES, NCONS, MINB, NPW>;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  const int n_tiles = (list ? n : L.nb) * (NC / KT);
  const int grid = n_tiles < MINB * sm_count() ? n_tiles : MINB * sm_count();
  OMG_LAUNCH_T(kern, grid, NCONS + 32 * NPW, smem, s, L, P, colour, list, n_tiles, fused);
  return true;
}
// shape of the ring kernel: OMG_RING_CFG = consumers * 1000 + stages * 100 + CTAs per SM * 10 + producer warps (tuning runs)
bool dispatch_smooth_ring(const LevelDev& L, const ParamsDev& P, int colour, const int* list, int n, int fused, cudaStream_t s) {
  static const int cfg = env_int("OMG_RING_CFG", 256222);
  switch (cfg) {
    case 256511: return launch_smooth_ring<16, 16, 5, 256, 1, 1>(L, P, colour, list, n, fused, s);
    case 256512: return launch_smooth_ring<16, 16, 5, 256, 1, 2>(L, P, colour, list, n, fused, s);
    case 512514: return launch_smooth_ring<16, 16, 5, 512, 1, 4>(L, P, colour, list, n, fused, s);
    case 512414: return launch_smooth_ring<16, 16, 4, 512, 1, 4>(L, P, colour, list, n, fused, s);
    case 256221: return launch_smooth_ring<16, 16, 2, 256, 2, 1>(L, P, colour, list, n, fused, s);
    case 256224: return launch_smooth_ring<16, 16, 2, 256, 2, 4>(L, P, colour, list, n, fused, s);
    case 256132: return launch_smooth_ring<16, 16, 1, 256, 3, 2>(L, P, colour, list, n, fused, s);
    case 256514: return launch_smooth_ring<16, 16, 5, 256, 1, 4>(L, P, colour, list, n, fused, s);
    case 512222: return launch_smooth_ring<16, 16, 2, 512, 2, 2>(L, P, colour, list, n, fused, s);
    default: return launch_smooth_ring<16, 16, 2, 256, 2, 2>(L, P, colour, list, n, fused, s);   // 0.270 ms per finest-level half-sweep at 512^3 (profiles/r02f_*)
  }
}

// =================================================================================================
// mg_smoother_gs: lexicographic Gauss-Seidel inside one box (smooth_boxes / box_gs_lpl, src/m_laplacian.f90:64-84 and
// m_helmholtz.f90:78-103 with the smoother type of m_multigrid.f90:380-410) against the ghost snapshot of
// omgk::gs_snapshot.  One CTA per box: phi and rhs arrive by two bulk copies (the colour-split box is contiguous), the four
// y / z ghost faces by a third, the two x ghosts of a column in registers.  Thread (j, k) owns the column of cells
// (0..NC-1, j, k) and updates cell i = w - j - k of hyperplane w; the cells of a hyperplane have the colour w & 1, their six
// neighbours the other one, and a lexicographic sweep only reads lower neighbours that are already new and upper ones that
// are still old: the hyperplane order performs the same arithmetic on the same values (as do_gs_sweep, omg_ops.cuh).
// Everything that depends on (j, k) alone -- which neighbour is a ghost, where it is stored -- is a per-thread constant
// (base pointer + stride of the other colour + shift of i), so a step has no data-dependent branch besides "is my cell in
// this hyperplane".  72 KB of shared memory: three boxes per SM.  Algorithmic traffic 24 B per cell + the snapshot.
// =================================================================================================
template <int NC>
struct GsSmem {
  static constexpr int HX = NC / 2;
  double Ph[2][NC][NC][HX];
  double R[2][NC][NC][HX];
  double G[4][NC][NC];       // ghost faces y-, y+, z-, z+ as do_gs_snapshot stores them: [face][t2][t1]
  uint64_t bar;
};
template <int NC>
__global__ void __launch_bounds__(kThreads, NC == 16 ? 3 : 8) k_gs_sweep_fast(const __grid_constant__ LevelDev L) {
  constexpr int HX = NC / 2, CS = NC * NC * HX, BOX = 2 * CS, PAIRS = NC * NC;
  OMG_DYN_SMEM(smem_raw, 128);
  GsSmem<NC>& S = *reinterpret_cast<GsSmem<NC>*>(smem_raw);
  const int tid = threadIdx.x, b = blockIdx.x;
  double* __restrict__ phi = L.var[0] + (size_t)b * L.bs;
  const double* __restrict__ gh = L.faces[0] + (size_t)b * 6 * PAIRS;
  if (tid == 0) {
    mbar_init(&S.bar, 1);
    mbar_arrive_expect_tx(&S.bar, (2 * BOX + 4 * PAIRS) * 8);
    bulk_g2s(&S.Ph[0][0][0][0], phi, BOX * 8, &S.bar);
    bulk_g2s(&S.R[0][0][0][0], L.var[1] + (size_t)b * L.bs, BOX * 8, &S.bar);
    bulk_g2s(&S.G[0][0][0], gh + 2 * PAIRS, 4 * PAIRS * 8, &S.bar);
  }
  const double idx2 = L.idr2[0], idy2 = L.idr2[1], idz2 = L.idr2[2], fac = L.fac;
  const bool act = tid < PAIRS;
  const int j = act ? tid % NC : 0, k = act ? tid / NC : 0, s0 = j + k;
  const double gx0 = act ? gh[k * NC + j] : 0.0, gx1 = act ? gh[PAIRS + k * NC + j] : 0.0;
  // neighbour columns: value(i, colour oc) = base[oc * os + (i >> sh)]
  double* const col = &S.Ph[0][k][j][0];
  const double* const rcol = &S.R[0][k][j][0];
  const double* const pyl = j > 0 ? col - HX : &S.G[0][k][0];
  const double* const pyh = j < NC - 1 ? col + HX : &S.G[1][k][0];
  const double* const pzl = k > 0 ? col - NC * HX : &S.G[2][j][0];
  const double* const pzh = k < NC - 1 ? col + NC * HX : &S.G[3][j][0];
  const int os_yl = j > 0 ? CS : 0, sh_yl = j > 0 ? 1 : 0;
  const int os_yh = j < NC - 1 ? CS : 0, sh_yh = j < NC - 1 ? 1 : 0;
  const int os_zl = k > 0 ? CS : 0, sh_zl = k > 0 ? 1 : 0;
  const int os_zh = k < NC - 1 ? CS : 0, sh_zh = k < NC - 1 ? 1 : 0;
  __syncthreads();
  mbar_wait(&S.bar, 0);
  for (int w = 0; w <= 3 * (NC - 1); ++w) {
    const int i = w - s0;
    if (act && i >= 0 && i < NC) {
      const int c = w & 1, oc = c ^ 1, m = i >> 1;
      const int il = (i > 0 ? i - 1 : 0) >> 1, ih = (i < NC - 1 ? i + 1 : NC - 1) >> 1;
      const double a_xl = col[oc * CS + il], a_xh = col[oc * CS + ih];
      const double ylo = pyl[oc * os_yl + (i >> sh_yl)], yhi = pyh[oc * os_yh + (i >> sh_yh)];
      const double zlo = pzl[oc * os_zl + (i >> sh_zl)], zhi = pzh[oc * os_zh + (i >> sh_zh)];
      const double r = rcol[c * CS + m];
      const double xlo = i > 0 ? a_xl : gx0, xhi = i < NC - 1 ? a_xh : gx1;
      col[c * CS + m] = fac * (idx2 * (xhi + xlo) + idy2 * (yhi + ylo) + idz2 * (zhi + zlo) - r);
    }
    __syncthreads();
  }
  fence_proxy_async();                // generic-proxy writes to Ph before the async proxy (TMA store) reads them
  __syncthreads();
  if (tid == 0) { bulk_s2g(phi, &S.Ph[0][0][0][0], BOX * 8); bulk_commit_wait_read(); }
}
template <int NC>
bool launch_gs_sweep(const LevelDev& L, cudaStream_t s) {
  static bool configured = false;
  const size_t smem = sizeof(GsSmem<NC>);
  auto kern = k_gs_sweep_fast<NC>;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  OMG_LAUNCH(kern, L.nb, smem, s, L);
  return true;
}

// launch with the programmatic-stream-serialization attribute: the kernel may start before its predecessor in the stream
// has finished and orders itself with pdl_sync()
bool g_pdl = false;
template <typename... KArgs, typename... Args>
bool launch_pdl(void (*kern)(KArgs...), int grid, size_t smem, cudaStream_t s, Args&&... args) {
#ifdef OMG_HOST_EMU
  return false;
#else
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...) == cudaSuccess;
#endif
}

template <int NC, int KT, int VC>
bool launch_smooth(const LevelDev& L, const ParamsDev& P, int colour, const int* list, int n, int fused, cudaStream_t s) {
  static bool configured = false;
  const size_t smem = sizeof(SmoothSmem<NC, KT, VC>);
  if (!configured) {
    if (cudaFuncSetAttribute(k_smooth_lpl_fast<NC, KT, VC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  const int grid = (list ? n : L.nb) * (NC / KT);
  if (g_pdl && !fused) return launch_pdl(k_smooth_lpl_fast<NC, KT, VC>, grid, smem, s, L, P, colour, list, fused, 1);
  auto kern = k_smooth_lpl_fast<NC, KT, VC>;
  OMG_LAUNCH(kern, grid, smem, s, L, P, colour, list, fused, 0);
  return true;
}


// =================================================================================================
// Both-colour staged stencil kernel, Laplacian / Helmholtz.  One thread per coarse cell = 2x2x2 fine cells
// (so the restriction sum needs no exchange and uses the reference's element order).
//   MODE_RESID_RESTRICT  residual_box + 2 x mg_restrict_lvl of update_coarse (src/m_multigrid.f90:355-363,
//                        m_restrict.f90:165-214): res = rhs - L(phi) stored, phi and res restricted onto the
//                        parent's octant.  26 B per fine cell (18 B + the stored fine residual).
//   MODE_RESID_MAX       max_residual_lvl (:296-311): res stored, max |res| -> atomic max.  24 B per cell.
//   MODE_CRHS            tail of update_coarse (:369-383) on parents: rhs = L(phi) + res, old = phi.  32 B.
// =================================================================================================
enum { MODE_RESID_RESTRICT = 0, MODE_RESID_MAX = 1, MODE_CRHS = 2 };

// VC = true (vlaplacian / vhelmholtz: box_vlpl src/m_vlaplacian.f90:131-189, box_vhelmh src/m_vhelmholtz.f90:144-205):
// eps is staged like phi (its ghost layer from the stored-face pool) and the operator becomes
// sum_f 2 idr2 a0 a_f / (a0 + a_f) (u_f - u0) [- lambda u0]; + 8 B per cell.
template <int NC, int KT>
struct BothSmem {
  static constexpr int HX = NC / 2;
  double Ph[2][KT][NC][HX];
  double zlo[2][NC][HX], zhi[2][NC][HX];
  double ylo[2][KT][HX], yhi[2][KT][HX];
  double xlo[KT][NC], xhi[KT][NC];
};
template <int NC, int KT, int NE>
struct StencilSmem : BothSmem<NC, KT> {
  typename std::conditional<(NE > 0), BothSmem<NC, KT>[(NE > 0 ? NE : 1)], NoSmem>::type E;   // NE coefficient arrays, see SmoothSmem
  double red[32];
  uint64_t bar;
};

// the six neighbours (face order -x, +x, -y, +y, -z, +z) of fine cell e = di | slot << 1 | dk << 2 of the 2x2x2 block
// (I, J, Kc) whose own values are u[8]; dj = slot ^ sw (see the row-order note in k_stencil_fast)
template <int NC, int KT>
__device__ __forceinline__ void block_neighbours(const BothSmem<NC, KT>& F, const double (&u)[8], int e, int I, int J, int Kc, int sw, double (&n)[6]) {
  constexpr int HC = NC / 2;
  const int di = e & 1, s = (e >> 1) & 1, dk = e >> 2;
  const int dj = s ^ sw, j = 2 * J + dj, kk = 2 * Kc + dk;
  const int oc = (di + dj + dk + 1) & 1;                       // colour of the six neighbours
  double xo;                                                   // x: the neighbour inside the block is u[e ^ 1]
  if (di == 0) xo = I > 0 ? F.Ph[oc][kk][j][I - 1] : F.xlo[kk][j];
  else xo = I < HC - 1 ? F.Ph[oc][kk][j][I + 1] : F.xhi[kk][j];
  n[0] = di ? u[e ^ 1] : xo; n[1] = di ? xo : u[e ^ 1];
  const int jo = dj ? j + 1 : j - 1;                           // y: inside neighbour u[e ^ 2]
  double yo;
  if (jo < 0) yo = F.ylo[oc][kk][I];
  else if (jo >= NC) yo = F.yhi[oc][kk][I];
  else yo = F.Ph[oc][kk][jo][I];
  n[2] = dj ? u[e ^ 2] : yo; n[3] = dj ? yo : u[e ^ 2];
  double zo;                                                   // z: inside neighbour u[e ^ 4]
  if (dk == 0) zo = kk > 0 ? F.Ph[oc][kk - 1][j][I] : F.zlo[oc][j][I];
  else zo = kk < KT - 1 ? F.Ph[oc][kk + 1][j][I] : F.zhi[oc][j][I];
  n[4] = dk ? u[e ^ 4] : zo; n[5] = dk ? zo : u[e ^ 4];
}

// STORE = false (option "store_fine_residual" = 0, for hosts that never read mg_ires): the residual of the level itself is
// not written -- nothing on the device reads it again (update_coarse consumes it in registers, the next residual pass
// overwrites it) -- which takes 8 B per cell off MODE_RESID_RESTRICT (26 -> 18 B) and MODE_RESID_MAX (24 -> 16 B).
template <int NC, int KT, int MODE, bool HELM, int MB, int NE, bool STORE>
__global__ void __launch_bounds__(kThreads, MB) k_stencil_fast(const __grid_constant__ LevelDev L, const __grid_constant__ LevelDev Lc,
                                                           const __grid_constant__ ParamsDev P, double* maxslot, int pdl) {
  constexpr int HX = NC / 2;
  constexpr int TILES = NC / KT;
  constexpr bool VC = NE > 0;
  constexpr int EY = NE == 3 ? 1 : 0, EZ = NE == 3 ? 2 : 0;      // which array serves the y / z faces (x: array 0)
  using SM = StencilSmem<NC, KT, NE>;
  OMG_DYN_SMEM(smem_raw, 128);
  SM& S = *reinterpret_cast<SM*>(smem_raw);
  const int tid = threadIdx.x;
  const int bq = blockIdx.x / TILES;
  const int b = (MODE == MODE_CRHS) ? L.parents_list[bq] : bq;
  const int k0 = (blockIdx.x % TILES) * KT;
  const double* __restrict__ phi = L.var[0];
  const size_t boxoff = (size_t)b * L.bs;

  // the tile itself depends on nothing but the box index: its copies start before the face table is read
  if (tid == 0) mbar_init(&S.bar, 1);
  pdl_sync(pdl);
  if (tid == 0) {
    mbar_expect_tx(&S.bar, (1 + NE) * 2 * KT * NC * HX * 8);
    if (TILES == 1) bulk_g2s(&S.Ph[0][0][0][0], phi + boxoff, 2 * KT * NC * HX * 8, &S.bar);   // whole box: one 32 KiB copy
    else {
      bulk_g2s(&S.Ph[0][0][0][0], phi + boxoff + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
      bulk_g2s(&S.Ph[1][0][0][0], phi + boxoff + L.cs + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
    }
    if constexpr (VC) {
#pragma unroll
      for (int ee = 0; ee < NE; ++ee) {
        const double* __restrict__ eps = L.var[4 + ee] + boxoff;
        if (TILES == 1) bulk_g2s(&S.E[ee].Ph[0][0][0][0], eps, 2 * KT * NC * HX * 8, &S.bar);
        else {
          bulk_g2s(&S.E[ee].Ph[0][0][0][0], eps + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
          bulk_g2s(&S.E[ee].Ph[1][0][0][0], eps + L.cs + (size_t)k0 * NC * HX, KT * NC * HX * 8, &S.bar);
        }
      }
    }
  }
  // rhs (or res) of this thread's first coarse cell: issued before the wait so that the loads overlap the bulk copies
  const double idx2 = L.idr2[0], idy2 = L.idr2[1], idz2 = L.idr2[2], lambda = P.lambda;
  const double* __restrict__ in1 = (MODE == MODE_CRHS ? L.var[3] : L.var[1]) + boxoff;   // res (crhs) or rhs
  double* __restrict__ out1 = (MODE == MODE_CRHS ? L.var[1] : L.var[3]) + boxoff;        // rhs (crhs) or res
  double* __restrict__ out2 = L.var[2] + boxoff;                                         // old (crhs)
  double vmax = 0.0;
  int p = -1, ox = 0, oy = 0, oz = 0;
  if (MODE == MODE_RESID_RESTRICT) {
    p = L.parent[b];
    const int off = L.coff[b];
    ox = off & 1023; oy = (off >> 10) & 1023; oz = (off >> 20) & 1023;
  }
  constexpr int HC = NC / 2, NCOARSE = (KT / 2) * HC * HC, NIT = NCOARSE / kThreads;
  static_assert(NCOARSE % kThreads == 0, "coarse cells must tile the CTA");
  const int cs = L.cs;
  // global offset of fine cell (di, slot s, dk) of coarse cell q (its colour is (di + dj + dk) & 1: k0, 2I, 2J, 2Kc are even)
  auto goff = [&](int q, int di, int s, int dk) -> size_t {
    const int I = q % HC, J = (q / HC) % HC, Kc = q / (HC * HC);
    const int dj = s ^ (J & 1);
    return (size_t)((di + dj + dk) & 1) * cs + ((size_t)(k0 + 2 * Kc + dk) * NC + (2 * J + dj)) * HX + I;
  };
  double nxt[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) nxt[e] = in1[goff(tid, e & 1, (e >> 1) & 1, e >> 2)];
  int kind[6], arg[6];
#pragma unroll
  for (int f = 0; f < 6; ++f) { kind[f] = L.fkind[b * 6 + f]; arg[f] = L.farg[b * 6 + f]; }
  const bool zlo_in = (k0 > 0), zhi_in = (k0 + KT < NC);
  __syncthreads();

  if (tid == 0) {
    uint32_t bytes = 0;
    if (zlo_in || kind[4] == FK_SAME) bytes += 2 * NC * HX * 8;
    if (zhi_in || kind[5] == FK_SAME) bytes += 2 * NC * HX * 8;
    if (VC && zlo_in) bytes += 2 * NC * HX * 8;
    if (VC && zhi_in) bytes += 2 * NC * HX * 8;
    mbar_arrive_expect_tx(&S.bar, bytes);
    if constexpr (VC) {
      const double* __restrict__ eps = L.var[4 + EZ] + boxoff;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (zlo_in) bulk_g2s(&S.E[EZ].zlo[c][0][0], eps + (size_t)c * L.cs + (size_t)(k0 - 1) * NC * HX, NC * HX * 8, &S.bar);
        if (zhi_in) bulk_g2s(&S.E[EZ].zhi[c][0][0], eps + (size_t)c * L.cs + (size_t)(k0 + KT) * NC * HX, NC * HX * 8, &S.bar);
      }
    }
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      if (zlo_in) bulk_g2s(&S.zlo[c][0][0], phi + boxoff + (size_t)c * L.cs + (size_t)(k0 - 1) * NC * HX, NC * HX * 8, &S.bar);
      else if (kind[4] == FK_SAME) bulk_g2s(&S.zlo[c][0][0], phi + (size_t)arg[4] * L.bs + (size_t)c * L.cs + (size_t)(NC - 1) * NC * HX, NC * HX * 8, &S.bar);
      if (zhi_in) bulk_g2s(&S.zhi[c][0][0], phi + boxoff + (size_t)c * L.cs + (size_t)(k0 + KT) * NC * HX, NC * HX * 8, &S.bar);
      else if (kind[5] == FK_SAME) bulk_g2s(&S.zhi[c][0][0], phi + (size_t)arg[5] * L.bs + (size_t)c * L.cs, NC * HX * 8, &S.bar);
    }
  }
  // y faces (rows j = NC-1 / 0 of the neighbours, both colours) and x faces (cells i = NC-1 / 0) by plain loads
  for (int r = tid; r < 2 * 2 * KT * HX; r += kThreads) {
    const int m = r % HX, kk = (r / HX) % KT, c = (r / (HX * KT)) & 1, hi = r / (2 * HX * KT);
    if (kind[2 + hi] == FK_SAME) {
      const double v = phi[(size_t)arg[2 + hi] * L.bs + (size_t)c * L.cs + ((size_t)(k0 + kk) * NC + (hi ? 0 : NC - 1)) * HX + m];
      if (hi) S.yhi[c][kk][m] = v; else S.ylo[c][kk][m] = v;
    }
  }
  for (int r = tid; r < 2 * KT * NC; r += kThreads) {
    const int j = r % NC, kk = (r / NC) % KT, hi = r / (NC * KT);
    if (kind[hi] == FK_SAME) {
      const int k = k0 + kk, i = hi ? 0 : NC - 1;
      const double v = phi[(size_t)arg[hi] * L.bs + (size_t)((i + j + k) & 1) * L.cs + ((size_t)k * NC + j) * HX + (i >> 1)];
      if (hi) S.xhi[kk][j] = v; else S.xlo[kk][j] = v;
    }
  }
  {
    // ghost planes materialised elsewhere (refinement boundaries, faces owned by another rank)
    const double* gpl = gpl_cur(L);
    if (!zlo_in && kind[4] == FK_PLANE)
      for (int r = tid; r < NC * NC; r += kThreads) { const int j = r / NC, i = r % NC; S.zlo[(i + j + 1) & 1][j][i >> 1] = ld_plane(gpl + (size_t)arg[4] * NC * NC + r); }
    if (!zhi_in && kind[5] == FK_PLANE)
      for (int r = tid; r < NC * NC; r += kThreads) { const int j = r / NC, i = r % NC; S.zhi[(i + j) & 1][j][i >> 1] = ld_plane(gpl + (size_t)arg[5] * NC * NC + r); }
    if (kind[2] == FK_PLANE)
      for (int r = tid; r < KT * NC; r += kThreads) { const int kk = r / NC, i = r % NC, k = k0 + kk; S.ylo[(i + k + 1) & 1][kk][i >> 1] = ld_plane(gpl + (size_t)arg[2] * NC * NC + k * NC + i); }
    if (kind[3] == FK_PLANE)
      for (int r = tid; r < KT * NC; r += kThreads) { const int kk = r / NC, i = r % NC, k = k0 + kk; S.yhi[(i + k) & 1][kk][i >> 1] = ld_plane(gpl + (size_t)arg[3] * NC * NC + k * NC + i); }
    if (kind[0] == FK_PLANE)
      for (int r = tid; r < KT * NC; r += kThreads) { const int kk = r / NC, j = r % NC; S.xlo[kk][j] = ld_plane(gpl + (size_t)arg[0] * NC * NC + (k0 + kk) * NC + j); }
    if (kind[1] == FK_PLANE)
      for (int r = tid; r < KT * NC; r += kThreads) { const int kk = r / NC, j = r % NC; S.xhi[kk][j] = ld_plane(gpl + (size_t)arg[1] * NC * NC + (k0 + kk) * NC + j); }
  }
  if constexpr (VC) {
    // ghost layer of eps: stored ghost faces (host data, any face kind), de-interleaved by colour
    const double* __restrict__ fc = L.faces[4] + (size_t)b * 6 * NC * NC;
    const double* __restrict__ fcy = L.faces[4 + EY] + (size_t)b * 6 * NC * NC;
    const double* __restrict__ fcz = L.faces[4 + EZ] + (size_t)b * 6 * NC * NC;
    if (!zlo_in)
      for (int r = tid; r < NC * NC; r += kThreads) { const int j = r / NC, i = r % NC; S.E[EZ].zlo[(i + j + 1) & 1][j][i >> 1] = fcz[4 * NC * NC + r]; }
    if (!zhi_in)
      for (int r = tid; r < NC * NC; r += kThreads) { const int j = r / NC, i = r % NC; S.E[EZ].zhi[(i + j) & 1][j][i >> 1] = fcz[5 * NC * NC + r]; }
    for (int r = tid; r < KT * NC; r += kThreads) {
      const int kk = r / NC, i = r % NC, k = k0 + kk;
      S.E[EY].ylo[(i + k + 1) & 1][kk][i >> 1] = fcy[2 * NC * NC + k * NC + i];
      S.E[EY].yhi[(i + k) & 1][kk][i >> 1] = fcy[3 * NC * NC + k * NC + i];
      S.E[0].xlo[kk][i] = fc[0 * NC * NC + k * NC + i];       // (here i runs over j)
      S.E[0].xhi[kk][i] = fc[1 * NC * NC + k * NC + i];
    }
  }
  mbar_wait(&S.bar, 0);
  __syncthreads();

  auto inner = [&](int i, int j, int kk) -> double { return S.Ph[(i + j + k0 + kk) & 1][kk][j][i >> 1]; };

  bool any_phys = false;
#pragma unroll
  for (int f = 0; f < 6; ++f) any_phys = any_phys || (kind[f] == FK_PHYS);
  if (any_phys) {
    for (int f = 0; f < 6; ++f) {
      if (kind[f] != FK_PHYS) continue;
      if (f == 4 && zlo_in) continue;
      if (f == 5 && zhi_in) continue;
      int type = L.ftype[b * 6 + f];
      if (type == 0) type = P.bc_type[f];
      double c0, c1, c2;
      bc_coeffs(type, L.dr[f >> 1], f, c0, c1, c2);
      const double* pl = arg[f] >= 0 ? L.bcpl + (size_t)arg[f] * NC * NC : nullptr;
      const double bconst = P.bc_value[f];
      const int n1 = (f & 1) ? NC - 1 : 0, n2 = (f & 1) ? NC - 2 : 1;
      if (f < 2) {
        for (int r = tid; r < KT * NC; r += kThreads) {
          const int kk = r / NC, j = r % NC;
          const double bv = pl ? pl[(k0 + kk) * NC + j] : bconst;
          const double g = c0 * bv + c1 * inner(n1, j, kk) + c2 * inner(n2, j, kk);
          if (f & 1) S.xhi[kk][j] = g; else S.xlo[kk][j] = g;
        }
      } else if (f < 4) {
        for (int r = tid; r < KT * NC; r += kThreads) {
          const int kk = r / NC, i = r % NC;
          const double bv = pl ? pl[(k0 + kk) * NC + i] : bconst;
          const double g = c0 * bv + c1 * inner(i, n1, kk) + c2 * inner(i, n2, kk);
          const int c = (i + (f & 1 ? NC : -1) + k0 + kk) & 1;
          if (f & 1) S.yhi[c][kk][i >> 1] = g; else S.ylo[c][kk][i >> 1] = g;
        }
      } else {
        const int kk1 = n1 - k0, kk2 = n2 - k0;
        for (int r = tid; r < NC * NC; r += kThreads) {
          const int j = r / NC, i = r % NC;
          const double bv = pl ? pl[j * NC + i] : bconst;
          const double g = c0 * bv + c1 * inner(i, j, kk1) + c2 * inner(i, j, kk2);
          const int c = (i + j + (f & 1 ? NC : -1)) & 1;
          if (f & 1) S.zhi[c][j][i >> 1] = g; else S.zlo[c][j][i >> 1] = g;
        }
      }
    }
    __syncthreads();
  }

  // ---- compute: one thread per coarse cell = 2x2x2 fine cells held in registers (8 + 24 shared-memory loads per
  // block instead of 8 x 7).  Rows j and j+2 are 16 doubles apart = the same banks, so odd-J threads visit their
  // two rows in the opposite order (slot s <-> dj = s ^ sw): a half-warp then reads rows 3 apart, conflict-free.
#pragma unroll
  for (int it = 0; it < NIT; ++it) {
    const int q = it * kThreads + tid;
    const int I = q % HC, J = (q / HC) % HC, Kc = q / (HC * HC);
    const int sw = J & 1;
    double cur[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) cur[e] = nxt[e];
    if (it + 1 < NIT) {
#pragma unroll
      for (int e = 0; e < 8; ++e) nxt[e] = in1[goff(q + kThreads, e & 1, (e >> 1) & 1, e >> 2)];
    }
    double u[8], r[8], ue[VC ? NE : 1][VC ? 8 : 1];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int di = e & 1, dj = ((e >> 1) & 1) ^ sw, dk = e >> 2;
      u[e] = S.Ph[(di + dj + dk) & 1][2 * Kc + dk][2 * J + dj][I];
      if constexpr (VC) {
#pragma unroll
        for (int ee = 0; ee < NE; ++ee) ue[ee][e] = S.E[ee].Ph[(di + dj + dk) & 1][2 * Kc + dk][2 * J + dj][I];
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int di = e & 1, s = (e >> 1) & 1, dk = e >> 2;
      const double u0 = u[e];
      double n[6];
      block_neighbours<NC, KT>(S, u, e, I, J, Kc, sw, n);
      double lap;
      if constexpr (VC) {
        double an[6], a0[6];
        if constexpr (NE == 3) {       // each array is only needed along its own direction (the rest is dead code)
          double ax[6], ay[6], az[6];
          block_neighbours<NC, KT>(S.E[0], ue[0], e, I, J, Kc, sw, ax);
          block_neighbours<NC, KT>(S.E[1], ue[1], e, I, J, Kc, sw, ay);
          block_neighbours<NC, KT>(S.E[2], ue[2], e, I, J, Kc, sw, az);
          an[0] = ax[0]; an[1] = ax[1]; an[2] = ay[2]; an[3] = ay[3]; an[4] = az[4]; an[5] = az[5];
#pragma unroll
          for (int f = 0; f < 6; ++f) a0[f] = ue[f >> 1][e];
        } else {
          block_neighbours<NC, KT>(S.E[0], ue[0], e, I, J, Kc, sw, an);
#pragma unroll
          for (int f = 0; f < 6; ++f) a0[f] = ue[0][e];
        }
        lap = P.coeff_math ? vlpl_value_tol(L.idr2, u0, n, a0, an) : vlpl_value(L.idr2, u0, n, a0, an);
      } else {
        lap = idx2 * (n[0] + n[1] - 2 * u0) + idy2 * (n[2] + n[3] - 2 * u0) + idz2 * (n[4] + n[5] - 2 * u0);
      }
      if (HELM) lap = lap - lambda * u0;
      const size_t g = goff(q, di, s, dk);
      if (MODE == MODE_CRHS) {
        out1[g] = lap + cur[e];
        out2[g] = u0;
      } else {
        r[e] = cur[e] - lap;
        if (STORE) out1[g] = r[e];
        if (MODE == MODE_RESID_MAX) vmax = fmax(vmax, fabs(r[e]));
      }
    }
    if (MODE == MODE_RESID_RESTRICT && p >= 0) {
      // 0.125 * sum(2x2x2) in Fortran element order (i fastest, then j, then k): undo the row swap
      double sphi = 0.0, sres = 0.0;
#pragma unroll
      for (int dk = 0; dk < 2; ++dk)
#pragma unroll
        for (int dj = 0; dj < 2; ++dj)
#pragma unroll
          for (int di = 0; di < 2; ++di) {
            const int ea = di | (dj << 1) | (dk << 2), eb = ea ^ 2;
            sphi = sphi + (sw ? u[eb] : u[ea]);
            sres = sres + (sw ? r[eb] : r[ea]);
          }
      const int ic = ox + I, jc = oy + J, kc = oz + (k0 >> 1) + Kc;
      const size_t gc = (size_t)p * Lc.bs + (size_t)((ic + jc + kc) & 1) * Lc.cs + ((size_t)kc * NC + jc) * HX + (ic >> 1);
      Lc.var[0][gc] = 0.125 * sphi;
      Lc.var[3][gc] = 0.125 * sres;
    }
  }
  if (MODE == MODE_RESID_MAX) {
    for (int o = 16; o > 0; o >>= 1) vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if ((tid & 31) == 0) S.red[tid >> 5] = vmax;
    __syncthreads();
    if (tid < 32) {
      vmax = tid < kThreads / 32 ? S.red[tid] : 0.0;
      for (int o = 16; o > 0; o >>= 1) vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
      if (tid == 0) atomicMax(reinterpret_cast<unsigned long long*>(maxslot), (unsigned long long)__double_as_longlong(vmax));
    }
  }
}

int g_store_res = 1;     // option "store_fine_residual"
template <int NC, int KT, int MODE, bool HELM, int MB, int VC, bool STORE>
bool launch_stencil_s(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  static bool configured = false;
  const size_t smem = sizeof(StencilSmem<NC, KT, VC>);
  if (!configured) {
    if (cudaFuncSetAttribute(k_stencil_fast<NC, KT, MODE, HELM, MB, VC, STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  if (g_pdl) return launch_pdl(k_stencil_fast<NC, KT, MODE, HELM, MB, VC, STORE>, nboxes * (NC / KT), smem, s, L, Lc, P, maxslot, 1);
  auto kern = k_stencil_fast<NC, KT, MODE, HELM, MB, VC, STORE>;
  OMG_LAUNCH(kern, nboxes * (NC / KT), smem, s, L, Lc, P, maxslot, 0);
  return true;
}
template <int NC, int KT, int MODE, bool HELM, int MB, int VC = 0>
bool launch_stencil(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  if (MODE != MODE_CRHS && !g_store_res) return launch_stencil_s<NC, KT, MODE, HELM, MB, VC, false>(L, Lc, P, nboxes, maxslot, s);
  return launch_stencil_s<NC, KT, MODE, HELM, MB, VC, true>(L, Lc, P, nboxes, maxslot, s);
}

template <int MODE, bool HELM>
bool dispatch_stencil_h(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  static const int mb = env_int("OMG_STENCIL_MB", 3);
  if (L.nc == 16) {
    if (mb == 2) return launch_stencil<16, 16, MODE, HELM, 2>(L, Lc, P, nboxes, maxslot, s);
    if (mb == 4) return launch_stencil<16, 16, MODE, HELM, 4>(L, Lc, P, nboxes, maxslot, s);
    return launch_stencil<16, 16, MODE, HELM, 3>(L, Lc, P, nboxes, maxslot, s);
  }
  if (L.nc == 32) return launch_stencil<32, 8, MODE, HELM, 2>(L, Lc, P, nboxes, maxslot, s);
  return false;
}
template <int MODE, bool HELM>
bool dispatch_stencil_vc(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  if (!L.var[4] || !L.faces[4]) return false;
  static const int vc_kt = env_int("OMG_VC_STENCIL_KT", 16);   // tuning: half-box tiles (8): 46 KB of shared memory instead of 91
  if (L.nc == 16) return vc_kt == 8 ? launch_stencil<16, 8, MODE, HELM, 3, 1>(L, Lc, P, nboxes, maxslot, s) : launch_stencil<16, 16, MODE, HELM, 2, 1>(L, Lc, P, nboxes, maxslot, s);
  if (L.nc == 32) return launch_stencil<32, 8, MODE, HELM, 1, 1>(L, Lc, P, nboxes, maxslot, s);
  return false;
}
// ahelmholtz (box_ahelmh src/m_ahelmholtz.f90:165-237): three coefficient arrays; half-box tiles at box 16 (4 x 24 KB)
template <int MODE>
bool dispatch_stencil_aniso(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  for (int e = 0; e < 3; ++e) if (!L.var[4 + e] || !L.faces[4 + e]) return false;
  static const int ah_kt = env_int("OMG_AH_STENCIL_KT", 8);
  if (L.nc == 16 && ah_kt == 16) return launch_stencil<16, 16, MODE, true, 1, 3>(L, Lc, P, nboxes, maxslot, s);
  if (L.nc == 16) return launch_stencil<16, 8, MODE, true, 2, 3>(L, Lc, P, nboxes, maxslot, s);
  if (L.nc == 32) return launch_stencil<32, 4, MODE, true, 1, 3>(L, Lc, P, nboxes, maxslot, s);
  return false;
}
template <int MODE>
bool dispatch_stencil(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, int nboxes, double* maxslot, cudaStream_t s) {
  if (P.op == OP_VLPL) return dispatch_stencil_vc<MODE, false>(L, Lc, P, nboxes, maxslot, s);
  if (P.op == OP_VHELMH) return dispatch_stencil_vc<MODE, true>(L, Lc, P, nboxes, maxslot, s);
  if (P.op == OP_AHELMH) return dispatch_stencil_aniso<MODE>(L, Lc, P, nboxes, maxslot, s);
  return P.op == OP_HELMH ? dispatch_stencil_h<MODE, true>(L, Lc, P, nboxes, maxslot, s)
                          : dispatch_stencil_h<MODE, false>(L, Lc, P, nboxes, maxslot, s);
}


// =================================================================================================
// correct_children fused with mg_prolong / mg_prolong_sparse (src/m_multigrid.f90:387-402,
// src/m_prolong.f90:124-156, :159-240), seen from the fine box: the (NC/2+2)^3 patch of the correction
// phi_c - old_c around the child's octant is staged in shared memory (ghost values through cell_at, so any
// face kind works), then every thread prolongs coarse cells onto 2x2x2 fine cells and adds them to phi.
//   18 B per fine cell (read + write phi_f, 1/8 of phi_c and old_c) + 1 B for the stored coarse res.
// =================================================================================================
template <int NC, int MB>
__global__ void __launch_bounds__(kThreads, MB) k_corr_prolong_fast(const __grid_constant__ LevelDev Lf, const __grid_constant__ LevelDev Lc,
                                                                const __grid_constant__ ParamsDev P, int pdl, int fused_put) {
  constexpr int HC = NC / 2, HX = NC / 2, PW = HC + 2, NIT = HC * HC * HC / kThreads;
  static_assert((HC * HC * HC) % kThreads == 0, "coarse cells must tile the CTA");
  OMG_DYN_SMEM(smem_raw, 16);
  double* corr = reinterpret_cast<double*>(smem_raw);       // [PW][PW][PW]
  const int tid = threadIdx.x, b = blockIdx.x;
  const int p = Lf.parent[b];
  const int off = Lf.coff[b];
  // in-kernel halo exchange (fused_put, as in k_smooth_lpl_fast): a box on a partition face stores its corrected face layers
  // into the neighbour ranks' plane pools; the GSRB half-sweep that follows waits for the peers' signal in its own
  // partition-face CTAs, so that no stand-alone ghost exchange (pack + put + signal + wait) runs between the two
  const int put_slot = fused_put ? Lf.put.slot_of[b] : -1;
  pdl_sync(pdl);
  if (p == -1) {                      // (uniform across the CTA)
    if (put_slot >= 0) put_faces<NC, NC>(Lf, put_slot, (size_t)b * Lf.bs, 0, 1, tid, fused_put);
    return;
  }
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  double* __restrict__ phi = Lf.var[0] + (size_t)b * Lf.bs;
  const int fcs = Lf.cs;
  // offset of fine cell e = di | dj << 1 | dk << 2 of coarse cell q inside the fine box
  auto goff = [&](int q, int e) -> int {
    const int I = q % HC, J = (q / HC) % HC, K = q / (HC * HC);
    const int di = e & 1, dj = (e >> 1) & 1, dk = e >> 2;
    return ((di + dj + dk) & 1) * fcs + ((2 * K + dk) * NC + (2 * J + dj)) * HX + I;
  };
  // fine phi of this thread's first coarse cell: in flight while the correction patch is gathered
  double nxt[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) nxt[e] = phi[goff(tid, e)];

  if (p <= -2) {   // parent on another rank: the patch arrived with the prolong exchange
    const double* __restrict__ patch = Lf.ppatch + (size_t)(-(p + 2)) * PW * PW * PW;
    for (int r = tid; r < PW * PW * PW; r += kThreads) corr[r] = patch[r];
  } else {
    // (a) the child's octant of the parent: always interior cells; res_c = phi_c - old_c is stored there (:396-398)
    const double* __restrict__ pc = Lc.var[0] + (size_t)p * Lc.bs;
    const double* __restrict__ po = Lc.var[2] + (size_t)p * Lc.bs;
    double* __restrict__ pr = Lc.var[3] + (size_t)p * Lc.bs;
    const int ccs = Lc.cs;
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int r = it * kThreads + tid;
      const int a = r % HC, bb = (r / HC) % HC, c = r / (HC * HC);
      const int ic = ox + a, jc = oy + bb, kc = oz + c;
      const int g = ((ic + jc + kc) & 1) * ccs + (kc * NC + jc) * HX + (ic >> 1);
      const double v = pc[g] - po[g];
      corr[((c + 1) * PW + (bb + 1)) * PW + (a + 1)] = v;
      pr[g] = v;
    }
    // (b) the six HC x HC halo faces of the patch: parent cells of the neighbouring octant or parent ghost cells
    // (edges and corners of the patch are never read by the 7-point prolongation)
    for (int r = tid; r < 6 * HC * HC; r += kThreads) {
      const int f = r / (HC * HC), t = r % (HC * HC), t1 = t % HC + 1, t2 = t / HC + 1;
      const int e = (f & 1) ? PW - 1 : 0;
      int a, bb, c;
      if (f < 2) { a = e; bb = t1; c = t2; } else if (f < 4) { a = t1; bb = e; c = t2; } else { a = t1; bb = t2; c = e; }
      const int ic = ox + a - 1, jc = oy + bb - 1, kc = oz + c - 1;
      corr[(c * PW + bb) * PW + a] = cell_at<GM_PHI>(Lc, P, Lc.var[0], nullptr, p, ic, jc, kc) - cell_at<GM_OLD>(Lc, P, Lc.var[2], nullptr, p, ic, jc, kc);
    }
  }
  __syncthreads();
#pragma unroll(NIT <= 4 ? NIT : 2)
  for (int it = 0; it < NIT; ++it) {
    const int q = it * kThreads + tid;
    double cur[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) cur[e] = nxt[e];
    if (it + 1 < NIT) {
#pragma unroll
      for (int e = 0; e < 8; ++e) nxt[e] = phi[goff(q + kThreads, e)];
    }
    const int I = q % HC, J = (q / HC) % HC, K = q / (HC * HC);
    const int c = ((K + 1) * PW + (J + 1)) * PW + (I + 1);
    const double f0 = 0.25 * corr[c];
    const double flx = 0.25 * corr[c - 1], fhx = 0.25 * corr[c + 1];
    const double fly = 0.25 * corr[c - PW], fhy = 0.25 * corr[c + PW];
    const double flz = 0.25 * corr[c - PW * PW], fhz = 0.25 * corr[c + PW * PW];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const double fine = f0 + ((e & 1) ? fhx : flx) + ((e & 2) ? fhy : fly) + ((e & 4) ? fhz : flz);
      phi[goff(q, e)] = cur[e] + fine;
    }
  }
  if (put_slot >= 0) put_faces<NC, NC>(Lf, put_slot, (size_t)b * Lf.bs, 0, 1, tid, fused_put);
}

// Box 16, TMA version: the child's 32 KiB of phi (both colours, contiguous) arrive by ONE bulk copy that is in flight
// from the first instruction -- it depends on nothing but the box index -- while the correction patch is gathered (a
// chain of dependent loads: parent index -> parent cells / face table -> neighbour cells); the update happens in place in
// shared memory and leaves by one bulk store.  The register-prefetch version above holds a CTA for ~8.9 us
// (4 CTAs / SM, 0.49 ms per launch on 32768 boxes = 0.75 of the measured HBM peak, profiles/r02d_*): its phi loads
// start only after the parent index has arrived and the second half of them only after the first half was consumed.
template <int NC, int MB>
__global__ void __launch_bounds__(kThreads, MB) k_corr_prolong_tma(const __grid_constant__ LevelDev Lf, const __grid_constant__ LevelDev Lc,
                                                               const __grid_constant__ ParamsDev P) {
  constexpr int HC = NC / 2, HX = NC / 2, PW = HC + 2, NIT = HC * HC * HC / kThreads, BOX = 2 * NC * NC * HX;
  static_assert((HC * HC * HC) % kThreads == 0, "coarse cells must tile the CTA");
  OMG_DYN_SMEM(smem_raw, 128);
  double* F = reinterpret_cast<double*>(smem_raw);                 // [2][NC][NC][HX]: the fine box as in global memory
  double* corr = F + BOX;                                          // [PW][PW][PW]
  uint64_t* bar = reinterpret_cast<uint64_t*>(corr + PW * PW * PW + (PW * PW * PW & 1));
  const int tid = threadIdx.x, b = blockIdx.x;
  double* __restrict__ phi = Lf.var[0] + (size_t)b * Lf.bs;
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_arrive_expect_tx(bar, BOX * 8);
    bulk_g2s(F, phi, BOX * 8, bar);
  }
  const int p = Lf.parent[b];
  const int off = Lf.coff[b];
  if (p == -1) {                      // no parent (uniform across the CTA): the copy must still land before the CTA may exit
    __syncthreads();
    mbar_wait(bar, 0);
    return;
  }
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  if (p <= -2) {   // parent on another rank: the patch arrived with the prolong exchange
    const double* __restrict__ patch = Lf.ppatch + (size_t)(-(p + 2)) * PW * PW * PW;
    for (int r = tid; r < PW * PW * PW; r += kThreads) corr[r] = patch[r];
  } else {
    const double* __restrict__ pc = Lc.var[0] + (size_t)p * Lc.bs;
    const double* __restrict__ po = Lc.var[2] + (size_t)p * Lc.bs;
    double* __restrict__ pr = Lc.var[3] + (size_t)p * Lc.bs;
    const int ccs = Lc.cs;
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int r = it * kThreads + tid;
      const int a = r % HC, bb = (r / HC) % HC, c = r / (HC * HC);
      const int ic = ox + a, jc = oy + bb, kc = oz + c;
      const int g = ((ic + jc + kc) & 1) * ccs + (kc * NC + jc) * HX + (ic >> 1);
      const double v = pc[g] - po[g];
      corr[((c + 1) * PW + (bb + 1)) * PW + (a + 1)] = v;
      pr[g] = v;
    }
    for (int r = tid; r < 6 * HC * HC; r += kThreads) {
      const int f = r / (HC * HC), t = r % (HC * HC), t1 = t % HC + 1, t2 = t / HC + 1;
      const int e = (f & 1) ? PW - 1 : 0;
      int a, bb, c;
      if (f < 2) { a = e; bb = t1; c = t2; } else if (f < 4) { a = t1; bb = e; c = t2; } else { a = t1; bb = t2; c = e; }
      const int ic = ox + a - 1, jc = oy + bb - 1, kc = oz + c - 1;
      corr[(c * PW + bb) * PW + a] = cell_at<GM_PHI>(Lc, P, Lc.var[0], nullptr, p, ic, jc, kc) - cell_at<GM_OLD>(Lc, P, Lc.var[2], nullptr, p, ic, jc, kc);
    }
  }
  __syncthreads();                    // corr complete (and the barrier initialisation visible to every thread)
  mbar_wait(bar, 0);
  // in place: one thread per coarse cell and iteration = 2x2x2 fine cells.  Rows 2J and 2J + 2 are 16 doubles apart = the
  // same banks, so odd-J threads visit their two rows in the opposite order (as in k_stencil_fast).
#pragma unroll
  for (int it = 0; it < NIT; ++it) {
    const int q = it * kThreads + tid;
    const int I = q % HC, J = (q / HC) % HC, K = q / (HC * HC);
    const int sw = J & 1;
    const int c = ((K + 1) * PW + (J + 1)) * PW + (I + 1);
    const double f0 = 0.25 * corr[c];
    const double flx = 0.25 * corr[c - 1], fhx = 0.25 * corr[c + 1];
    const double fly = 0.25 * corr[c - PW], fhy = 0.25 * corr[c + PW];
    const double flz = 0.25 * corr[c - PW * PW], fhz = 0.25 * corr[c + PW * PW];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int di = e & 1, dj = ((e >> 1) & 1) ^ sw, dk = e >> 2;
      const double fine = f0 + (di ? fhx : flx) + (dj ? fhy : fly) + (dk ? fhz : flz);
      const int g = ((di + dj + dk) & 1) * (NC * NC * HX) + ((2 * K + dk) * NC + (2 * J + dj)) * HX + I;
      F[g] = F[g] + fine;
    }
  }
  fence_proxy_async();                // generic-proxy writes to F before the async proxy (TMA store) reads them
  __syncthreads();
  if (tid == 0) { bulk_s2g(phi, F, BOX * 8); bulk_commit_wait_read(); }
}

template <int NC, int MB>
bool launch_corr_prolong_tma(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, cudaStream_t s) {
  static bool configured = false;
  constexpr int PW = NC / 2 + 2;
  const size_t smem = sizeof(double) * (2 * NC * NC * (NC / 2) + PW * PW * PW + (PW * PW * PW & 1)) + 16;
  auto kern = k_corr_prolong_tma<NC, MB>;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  OMG_LAUNCH(kern, Lf.nb, smem, s, Lf, Lc, P);
  return true;
}

template <int NC, int MB>
bool launch_corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int fused, cudaStream_t s) {
  static bool configured = false;
  const size_t smem = sizeof(double) * (NC / 2 + 2) * (NC / 2 + 2) * (NC / 2 + 2);
  if (!configured) {
    if (cudaFuncSetAttribute(k_corr_prolong_fast<NC, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
    configured = true;
  }
  if (g_pdl && !fused) return launch_pdl(k_corr_prolong_fast<NC, MB>, Lf.nb, smem, s, Lf, Lc, P, 1, 0);
  auto kern = k_corr_prolong_fast<NC, MB>;
  OMG_LAUNCH(kern, Lf.nb, smem, s, Lf, Lc, P, 0, fused);
  return true;
}

}  // namespace

namespace omgf {

void set_pdl(bool on) { g_pdl = on; }
void set_ring(int mode) { g_ring = mode; }
void set_store_res(int on) { g_store_res = on; }

bool smooth_applies(const LevelDev& L, const ParamsDev& P) {
  if (L.nc != 16 && L.nc != 32) return false;
  if (P.op == OP_VLPL || P.op == OP_VHELMH) return L.var[4] && L.faces[4];
  if (P.op == OP_AHELMH) return L.var[4] && L.faces[4] && L.var[5] && L.faces[5] && L.var[6] && L.faces[6];
  return P.op == OP_LPL || P.op == OP_HELMH;
}
// will smooth() run the persistent ring kernel on n boxes of this level?
bool smooth_uses_ring(const LevelDev& L, const ParamsDev& P, int n) {
  return L.nc == 16 && (P.op == OP_LPL || P.op == OP_HELMH) && g_ring && (n >= 4 * sm_count() || g_ring == 2);
}
bool smooth(const LevelDev& L, const ParamsDev& P, int colour, bool all_fast, const int* list, int n, int fused, cudaStream_t s) {
  if (!all_fast || L.nb == 0) return false;
  if (list && n == 0) return true;
  if (P.op == OP_VLPL || P.op == OP_VHELMH) {
    if (!L.var[4] || !L.faces[4]) return false;
    static const int vc_kt = env_int("OMG_VC_KT", 16);         // tuning: half-box tiles (8) give 5 CTAs per SM instead of 3
    if (L.nc == 16) return vc_kt == 8 ? launch_smooth<16, 8, 1>(L, P, colour, list, n, fused, s) : launch_smooth<16, 16, 1>(L, P, colour, list, n, fused, s);
    if (L.nc == 32) return launch_smooth<32, 8, 1>(L, P, colour, list, n, fused, s);
    return false;
  }
  if (P.op == OP_AHELMH) {
    for (int e = 0; e < 3; ++e) if (!L.var[4 + e] || !L.faces[4 + e]) return false;
    static const int ah_kt = env_int("OMG_AH_KT", 16);       // whole-box tiles: 1.06 vs 1.25 ms per half-sweep (profiles/r02s_summary.txt)
    if (L.nc == 16 && ah_kt == 16) return launch_smooth<16, 16, 3>(L, P, colour, list, n, fused, s);
    if (L.nc == 16) return launch_smooth<16, 8, 3>(L, P, colour, list, n, fused, s);     // half-box tiles: 4 x 12 KB of shared memory
    if (L.nc == 32) return launch_smooth<32, 8, 3>(L, P, colour, list, n, fused, s);
    return false;
  }
  if (P.op != OP_LPL && P.op != OP_HELMH) return false;
  if (smooth_uses_ring(L, P, list ? n : L.nb)) return dispatch_smooth_ring(L, P, colour, list, n, fused, s);
  if (L.nc == 16) return launch_smooth<16, 16, false>(L, P, colour, list, n, fused, s);
  if (L.nc == 32) return launch_smooth<32, 8, false>(L, P, colour, list, n, fused, s);
  return false;
}
bool gs_sweep(const LevelDev& L, const ParamsDev& P, cudaStream_t s) {
  static const int on = env_int("OMG_GS_FAST", 1);          // 0: the general hyperplane sweep of omg_kernels.cu (A/B runs)
  if (!on || L.nb == 0 || !L.faces[0] || (P.op != OP_LPL && P.op != OP_HELMH)) return false;
  if (L.nc == 16) return launch_gs_sweep<16>(L, s);
  if (L.nc == 8) return launch_gs_sweep<8>(L, s);
  return false;
}
bool resid_restrict(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, bool all_fast, cudaStream_t s) {
  if (!all_fast || Lf.nb == 0 || Lc.nc != Lf.nc) return false;
  return dispatch_stencil<MODE_RESID_RESTRICT>(Lf, Lc, P, Lf.nb, nullptr, s);
}
bool resid_max(const LevelDev& L, const ParamsDev& P, bool all_fast, double* maxslot, cudaStream_t s) {
  if (!all_fast || L.nb == 0 || !maxslot) return false;
  return dispatch_stencil<MODE_RESID_MAX>(L, L, P, L.nb, maxslot, s);
}
bool crhs(const LevelDev& L, const ParamsDev& P, bool all_fast, cudaStream_t s) {
  if (!all_fast || L.n_parents == 0) return false;
  return dispatch_stencil<MODE_CRHS>(L, L, P, L.n_parents, nullptr, s);
}
bool corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, bool, int fused, cudaStream_t s) {
  if (Lf.nb == 0 || Lc.nc != Lf.nc) return false;
  static const int mb = env_int("OMG_PROLONG_MB", 4);
  static const int tma = env_int("OMG_PROLONG_TMA", 0);        // CTAs per SM of the TMA version; 0 (default): register-prefetch version --
                                                               // measured equal within 1 % (0.488 / 0.492 / 0.492 / 0.516 ms for 0 / 3 / 4 / 5, profiles/r02f_*)
  // fused = 1: partition-face boxes put their corrected faces into the peers' plane pools (register-prefetch version only)
  if (Lf.nc == 16 && tma && !g_pdl && !fused) return tma == 4 ? launch_corr_prolong_tma<16, 4>(Lf, Lc, P, s) : tma == 3 ? launch_corr_prolong_tma<16, 3>(Lf, Lc, P, s) : launch_corr_prolong_tma<16, 5>(Lf, Lc, P, s);
  if (Lf.nc == 16) return mb == 4 ? launch_corr_prolong<16, 4>(Lf, Lc, P, fused, s) : mb == 2 ? launch_corr_prolong<16, 2>(Lf, Lc, P, fused, s) : launch_corr_prolong<16, 3>(Lf, Lc, P, fused, s);
  if (Lf.nc == 32) return launch_corr_prolong<32, 3>(Lf, Lc, P, fused, s);
  return false;
}

}  // namespace omgf
