// CTA emulator: a host build of the TMA-staged kernels (octree-mg_b200/csrc/omg_fast.cu) for CPU tests.
//
// Test infrastructure.  g++ compiles the SAME kernel source the GPU runs -- omg_fast.cu is included below -- with the
// CUDA execution model reduced to what these kernels use: the 256 threads of a CTA are 256 host threads, __syncthreads is
// a barrier, warp shuffles go through a per-warp exchange buffer, a TMA bulk copy (cp.async.bulk + mbarrier) is a memcpy
// whose alignment, size, destination range and expect_tx accounting are CHECKED (a mismatch there would hang or fault
// on the GPU), CTAs run one after the other.  What this proves: indexing, staging, boundary handling and arithmetic of the
// fast kernels against the general per-cell path (omg_ops.cuh do_*, the path the GPU parity tests and
// test_device_host.py pin to the oracle), bit for bit.  What it cannot show: races between CTAs, performance.
#include <pthread.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#define OMG_EMU_THREADS 1
#define OMG_HOST_EMU 1

struct FakeDim { unsigned x = 0, y = 0, z = 0; };
namespace emu {
constexpr int MAXT = 1024;
unsigned char* g_smem = nullptr;      // the CTA's dynamic shared memory
size_t g_smem_bytes = 0;              // as passed to the launch
pthread_barrier_t g_cta_bar, g_warp_bar[MAXT / 32], g_named_bar;
int g_named_count = 256;
double g_shfl[MAXT];
int g_shfl_i[MAXT];
// mbarriers with their real semantics (the persistent ring kernel runs a producer warp ahead of the consumers): pending
// arrivals + pending transaction bytes; a phase completes when both reach zero.  One lock for all of them.
struct MBar { int count = 0, pending = 0; long long tx = 0; unsigned phase = 0; };
std::mutex g_mb_m;
std::condition_variable g_mb_cv;
std::map<const void*, MBar> g_mb;
inline void mb_check(MBar& b) { if (b.pending == 0 && b.tx == 0) { b.phase++; b.pending = b.count; g_mb_cv.notify_all(); } }
std::atomic<long long> g_tx_expected{0}, g_tx_done{0};
std::atomic<int> g_violations{0};
void violation(const char* what) {
  if (g_violations.fetch_add(1) < 5) std::fprintf(stderr, "fast_emu: %s\n", what);
}
}  // namespace emu

static thread_local FakeDim threadIdx, blockIdx;
static FakeDim blockDim = {256, 1, 1}, gridDim = {1, 1, 1};
static inline void __syncthreads() { pthread_barrier_wait(&emu::g_cta_bar); }
static inline void __syncwarp() { pthread_barrier_wait(&emu::g_warp_bar[threadIdx.x >> 5]); }   // every use is warp-uniform
static inline void __threadfence_system() { std::atomic_thread_fence(std::memory_order_seq_cst); }
static inline double __shfl_xor_sync(unsigned, double v, int o) {
  const int t = (int)threadIdx.x;
  emu::g_shfl[t] = v;
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  const double r = emu::g_shfl[t ^ o];
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  return r;
}
static inline double __shfl_sync(unsigned, double v, int src) {
  const int t = (int)threadIdx.x;
  emu::g_shfl[t] = v;
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  const double r = emu::g_shfl[(t & ~31) | (src & 31)];
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  return r;
}
static inline unsigned long long atomicMax(unsigned long long* a, unsigned long long v) {
  unsigned long long o = __atomic_load_n(a, __ATOMIC_RELAXED);
  while (o < v && !__atomic_compare_exchange_n(a, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_RELAXED)) {}
  return o;
}
static inline unsigned int atomicAdd(unsigned int* a, unsigned int v) { return __atomic_fetch_add(a, v, __ATOMIC_SEQ_CST); }
static inline int atomicExch(int* a, int v) { return __atomic_exchange_n(a, v, __ATOMIC_SEQ_CST); }

#include "harness.cpp"   // HostLevel, build_level, randomise, params + the general device path (omg_ops.cuh)

// ---- stand-ins for the PTX helper block of omg_fast.cu ------------------------------------------------------
static inline bool in_smem(const void* p, size_t bytes) {
  const unsigned char* d = static_cast<const unsigned char*>(p);
  return d >= emu::g_smem && d + bytes <= emu::g_smem + emu::g_smem_bytes;
}
static inline emu::MBar& mb_of(uint64_t* bar) {
  if (!in_smem(bar, 8) || ((uintptr_t)bar & 7)) emu::violation("mbarrier outside the CTA's shared memory or misaligned");
  auto it = emu::g_mb.find(bar);
  if (it == emu::g_mb.end()) { emu::violation("mbarrier used before mbarrier.init"); return emu::g_mb[bar]; }
  return it->second;
}
static inline void mbar_init(uint64_t* bar, uint32_t count) {
  std::lock_guard<std::mutex> lk(emu::g_mb_m);
  emu::MBar b; b.count = b.pending = (int)count;
  emu::g_mb[bar] = b;
}
static inline void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  std::lock_guard<std::mutex> lk(emu::g_mb_m);
  mb_of(bar).tx += bytes; emu::g_tx_expected += bytes;
}
static inline void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  std::lock_guard<std::mutex> lk(emu::g_mb_m);
  emu::MBar& b = mb_of(bar);
  b.tx += bytes; emu::g_tx_expected += bytes;
  if (b.pending <= 0) emu::violation("mbarrier: more arrivals than the init count");
  b.pending--; emu::mb_check(b);
}
static inline void mbar_arrive(uint64_t* bar) {
  std::lock_guard<std::mutex> lk(emu::g_mb_m);
  emu::MBar& b = mb_of(bar);
  if (b.pending <= 0) emu::violation("mbarrier: more arrivals than the init count");
  b.pending--; emu::mb_check(b);
}
// try_wait.parity: true once the phase of that parity has completed, i.e. the phase in progress has the other parity
static inline void mbar_wait(uint64_t* bar, uint32_t parity) {
  std::unique_lock<std::mutex> lk(emu::g_mb_m);
  emu::MBar& b = mb_of(bar);
  if (!emu::g_mb_cv.wait_for(lk, std::chrono::seconds(60), [&] { return (b.phase & 1u) != (parity & 1u); }))
    emu::violation("mbarrier wait: the phase never completed (the GPU would hang)");
}
static inline void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  if (((uintptr_t)dst & 15) || ((uintptr_t)src & 15) || (bytes & 15) || bytes == 0) emu::violation("bulk copy: address or size not a multiple of 16");
  if (!in_smem(dst, bytes)) emu::violation("bulk copy: destination outside the CTA's shared memory");
  std::memcpy(dst, src, bytes);
  std::lock_guard<std::mutex> lk(emu::g_mb_m);
  emu::MBar& b = mb_of(bar);
  b.tx -= bytes; emu::g_tx_done += bytes; emu::mb_check(b);
}
// TMA store: the source must lie in the CTA's shared memory; alignment rules as for the load
static inline void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  if (((uintptr_t)dst & 15) || ((uintptr_t)src & 15) || (bytes & 15) || bytes == 0) emu::violation("bulk store: address or size not a multiple of 16");
  if (!in_smem(src, bytes)) emu::violation("bulk store: source outside the CTA's shared memory");
  std::memcpy(dst, src, bytes);
}
static inline void bulk_commit_wait_read() {}
static inline void cp_async8(void* dst, const void* src) {
  if (((uintptr_t)dst & 7) || ((uintptr_t)src & 7)) emu::violation("cp.async: address not a multiple of 8");
  if (!in_smem(dst, 8)) emu::violation("cp.async: destination outside the CTA's shared memory");
  std::memcpy(dst, src, 8);
}
static inline void cp_async16(void* dst, const void* src) {
  if (((uintptr_t)dst & 15) || ((uintptr_t)src & 15)) emu::violation("cp.async: address not a multiple of 16");
  if (!in_smem(dst, 16)) emu::violation("cp.async: destination outside the CTA's shared memory");
  std::memcpy(dst, src, 16);
}
static inline void fence_proxy_async() {}
static inline void cp_async_arrive(uint64_t* bar) { mbar_arrive(bar); }
static inline void consumer_sync(int nthreads) {
  if (nthreads != emu::g_named_count) emu::violation("named barrier: thread count differs from the CTA's consumer threads");
  pthread_barrier_wait(&emu::g_named_bar);
}
static inline int shfl_int(int v, int src) {
  const int t = (int)threadIdx.x;
  emu::g_shfl_i[t] = v;
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  const int r = emu::g_shfl_i[(t & ~31) | (src & 31)];
  pthread_barrier_wait(&emu::g_warp_bar[t >> 5]);
  return r;
}
// the emulated ranks run one after the other, so a flag a kernel would wait for must already be there
static inline bool wait_flag(const unsigned long long* f, unsigned long long want, int*, unsigned long long, int) {
  if (__atomic_load_n(f, __ATOMIC_ACQUIRE) < want) emu::violation("wait_flag: the peer's signal is missing (the GPU would time out)");
  return true;
}
static inline void signal_flag(unsigned long long* f, unsigned long long v) { __atomic_store_n(f, v, __ATOMIC_RELEASE); }
static inline double ld_plane(const double* p) { return *p; }
#define OMG_DYN_SMEM(name, al) unsigned char* name = emu::g_smem
#define OMG_PDL_ASM() do {} while (0)
#define OMG_LAUNCH(kern, grid, smem, stream, ...) emu::launch((grid), 256, (smem), [&]() { kern(__VA_ARGS__); })
#define OMG_LAUNCH_T(kern, grid, threads, smem, stream, ...) emu::launch((grid), (threads), (smem), [&]() { kern(__VA_ARGS__); })
#define OMG_TAIL_DYN_SMEM(name) double* name = reinterpret_cast<double*>(emu::g_smem)
#define cudaFuncSetAttribute(...) cudaSuccess
#undef __launch_bounds__
#define __launch_bounds__(...)
#undef __grid_constant__
#define __grid_constant__

namespace emu {
void launch(int grid, int nthreads, size_t smem, const std::function<void()>& body) {
  static unsigned char* buf = static_cast<unsigned char*>(aligned_alloc(1024, 256 * 1024));
  g_smem = buf; g_smem_bytes = smem;
  if (smem > 227 * 1024) violation("launch: more than 227 KB of shared memory");
  blockDim.x = (unsigned)nthreads; gridDim.x = (unsigned)grid; if (getenv("FE_TRACE")) std::fprintf(stderr, "launch grid %d threads %d smem %zu\n", grid, nthreads, smem);
  pthread_barrier_init(&g_cta_bar, nullptr, nthreads);
  g_named_count = nthreads > 256 ? (nthreads >= 512 ? 512 : 256) : 256;      // ring kernels: 256 or 512 consumer threads + producer warps
  pthread_barrier_init(&g_named_bar, nullptr, g_named_count);
  for (auto& w : g_warp_bar) pthread_barrier_init(&w, nullptr, 32);
  std::vector<std::thread> th;
  th.reserve(nthreads);
  for (int t = 0; t < nthreads; ++t)
    th.emplace_back([&, t]() {
      for (int b = 0; b < grid; ++b) {
        threadIdx.x = (unsigned)t; blockIdx.x = (unsigned)b;
        if (t == 0) { g_tx_expected = 0; g_tx_done = 0; std::memset(g_smem, 0xff, smem); g_mb.clear(); }   // NaN-fill: stale reads show
        pthread_barrier_wait(&g_cta_bar);
        body();
        pthread_barrier_wait(&g_cta_bar);
        if (t == 0 && g_tx_expected != g_tx_done) violation("mbarrier: expected bytes differ from the bytes copied (the GPU would hang)");
      }
    });
  for (auto& x : th) x.join();
  pthread_barrier_destroy(&g_cta_bar);
  pthread_barrier_destroy(&g_named_bar);
  for (auto& w : g_warp_bar) pthread_barrier_destroy(&w);
}
}  // namespace emu

#include "../../octree-mg_b200/csrc/omg_fast.cu"

// the tail interpreter: static __shared__ arrays become function-level statics (one CTA at a time)
#undef __shared__
#define __shared__ static
#include "../../octree-mg_b200/csrc/omg_tail.cuh"

namespace {

void copy_data(HostLevel& dst, const HostLevel& src) {   // same shape: values only (the LevelDev pointers stay the level's own)
  for (int v = 0; v < 4; ++v) std::copy(src.var[v].begin(), src.var[v].end(), dst.var[v].begin());
  for (int d = 0; d < 3; ++d) {
    std::copy(src.eps[d].begin(), src.eps[d].end(), dst.eps[d].begin());
    std::copy(src.epsf[d].begin(), src.epsf[d].end(), dst.epsf[d].begin());
  }
}
void set_fac(HostLevel& H, int op, double lambda) {
  const double sum = H.L.idr2[0] + H.L.idr2[1] + H.L.idr2[2];
  H.L.fac = (op == OP_HELMH) ? 1.0 / (2 * sum + lambda) : 0.5 / sum;
}
void link_children(HostLevel& F, int ncf, int nbx, int nby) {   // fine boxes = octants of the coarse boxes, x fastest
  for (int b = 0; b < F.L.nb; ++b) {
    const int fx = b % (2 * nbx), fy = (b / (2 * nbx)) % (2 * nby), fz = b / (4 * nbx * nby);
    F.parent[b] = ((fz / 2) * nby + fy / 2) * nbx + fx / 2;
    const int h = ncf / 2;
    F.coff[b] = ((fx & 1) * h) | (((fy & 1) * h) << 10) | (((fz & 1) * h) << 20);
  }
}

template <int OP>
int one_level(int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda, uint64_t seed) {
  HostLevel G, F;                    // general path / fast kernels
  build_level(G, nc, nbx, nby, nbz, bc_types, 0.01);
  build_level(F, nc, nbx, nby, nbz, bc_types, 0.01);
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  set_fac(G, OP, lambda); set_fac(F, OP, lambda);
  randomise(G, seed);
  copy_data(F, G);
  std::vector<int> all(G.L.nb);
  for (int b = 0; b < G.L.nb; ++b) all[b] = b;
  G.L.parents_list = F.L.parents_list = all.data(); G.L.n_parents = F.L.n_parents = G.L.nb;
  int bad = 0;
  for (int n = 0; n < 4; ++n) {      // two GSRB sweeps
    for (int b = 0; b < G.L.nb; ++b) do_smooth<OP>(G.L, P, b, n & 1, 0, 1);
    if (!omgf::smooth(F.L, P, n & 1, true, nullptr, 0, 0, nullptr)) return -1;
    bad += differing(G, F) ? 1 : 0;
  }
  {                                  // a box list (the partition-face / interior split launches subsets)
    std::vector<int> some;
    for (int b = 0; b < G.L.nb; b += 2) some.push_back(b);
    for (int b : some) do_smooth<OP>(G.L, P, b, 0, 0, 1);
    if (!omgf::smooth(F.L, P, 0, true, some.data(), (int)some.size(), 0, nullptr)) return -1;
    bad += differing(G, F) ? 1 : 0;
  }
  double mg = 0.0, mf = 0.0;         // residual and its maximum
  for (int b = 0; b < G.L.nb; ++b) mg = fmax(mg, do_resid<OP>(G.L, P, b, 0, 1));
  if (!omgf::resid_max(F.L, P, true, &mf, nullptr)) return -1;
  bad += (std::memcmp(&mg, &mf, 8) != 0) + (differing(G, F) ? 1 : 0);
  for (int b = 0; b < G.L.nb; ++b) do_crhs<OP>(G.L, P, b, 0, 1);   // coarse right-hand side on every box
  if (!omgf::crhs(F.L, P, true, nullptr)) return -1;
  bad += differing(G, F) ? 1 : 0;
  return bad;
}

template <int OP>
int two_levels(int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda, uint64_t seed) {
  HostLevel Gf, Gc, Ff, Fc;
  build_level(Gf, nc, 2 * nbx, 2 * nby, 2 * nbz, bc_types, 0.01); build_level(Gc, nc, nbx, nby, nbz, bc_types, 0.02);
  build_level(Ff, nc, 2 * nbx, 2 * nby, 2 * nbz, bc_types, 0.01); build_level(Fc, nc, nbx, nby, nbz, bc_types, 0.02);
  link_children(Gf, nc, nbx, nby); link_children(Ff, nc, nbx, nby);
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  for (HostLevel* H : {&Gf, &Gc, &Ff, &Fc}) set_fac(*H, OP, lambda);
  randomise(Gf, seed); randomise(Gc, seed + 17);
  copy_data(Ff, Gf); copy_data(Fc, Gc);
  int bad = 0;
  // update_coarse without its coarse right-hand side: residual + restriction of phi and res
  for (int b = 0; b < Gf.L.nb; ++b) { do_resid<OP>(Gf.L, P, b, 0, 1); do_restrict(Gf.L, Gc.L, 0, b, 0, 1); do_restrict(Gf.L, Gc.L, 3, b, 0, 1); }
  if (!omgf::resid_restrict(Ff.L, Fc.L, P, true, nullptr)) return -1;
  bad += (differing(Gf, Ff) ? 1 : 0) + (differing(Gc, Fc) ? 1 : 0);
  // correct_children
  randomise(Gc, seed + 31); copy_data(Fc, Gc);
  for (int b = 0; b < Gf.L.nb; ++b) do_corr_prolong(Gf.L, Gc.L, P, b, 0, 1);
  if (!omgf::corr_prolong(Ff.L, Fc.L, P, true, 0, nullptr)) return -1;
  bad += (differing(Gf, Ff) ? 1 : 0) + (differing(Gc, Fc) ? 1 : 0);
  return bad;
}


// ---- the in-kernel halo exchange (PutDev) between emulated ranks -------------------------------------------------------
// A brick of boxes is cut into n_ranks contiguous chunks of the x-fastest box order.  Every rank gets its own level
// (faces towards another rank are FK_PLANE entries of its double-buffered plane pool), the PutDev tables the library
// builds in sync_meta (omg_core.cu) and plain pointers to the other ranks' pools and flags where the GPU has peer
// mappings.  The ranks run each half-sweep one after the other; the state must follow the single-rank general path.
struct RankLevel {
  HostLevel H;
  std::vector<int> gid, slot_of, ptr, item, launch;
  std::vector<double> pool;
  std::vector<unsigned long long> flags;
  unsigned long long epoch = 0;
  unsigned int done = 0;
  int err = 0, n_planes = 0;
  std::vector<double*> peer_base; std::vector<long long> peer_stride; std::vector<unsigned long long*> peer_flags;
};

template <int OP>
int ranks_case(int nc, int nbx, int nby, int nbz, int n_ranks, int bnd_first, const int* bc_types, const double* bc_values,
               double lambda, uint64_t seed) {
  const int nb = nbx * nby * nbz;
  HostLevel G;
  build_level(G, nc, nbx, nby, nbz, bc_types, 0.01);
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  set_fac(G, OP, lambda);
  randomise(G, seed);
  auto owner = [&](int b) { return (int)((long long)b * n_ranks / nb); };
  std::vector<RankLevel> R(n_ranks);
  std::vector<int> local_of(nb);
  for (int b = 0; b < nb; ++b) { RankLevel& r = R[owner(b)]; local_of[b] = (int)r.gid.size(); r.gid.push_back(b); }
  struct Send { int box, f, to, plane; };
  std::vector<std::vector<Send>> sends(n_ranks);
  const int save_planes = g_planes; g_planes = 0;
  for (int me = 0; me < n_ranks; ++me) {
    RankLevel& r = R[me];
    const int nl = (int)r.gid.size();
    if (nl == 0) { g_planes = save_planes; return -2; }
    build_level(r.H, nc, nl, 1, 1, bc_types, 0.01);
    set_fac(r.H, OP, lambda);
    for (int lb = 0; lb < nl; ++lb) {
      const int g = r.gid[lb];
      const int pos[3] = {g % nbx, (g / nbx) % nby, g / (nbx * nby)}, ext[3] = {nbx, nby, nbz}, stride[3] = {1, nbx, nbx * nby};
      for (int f = 0; f < 6; ++f) {
        const int d = f >> 1, q = pos[d] + ((f & 1) ? 1 : -1);
        int& kind = r.H.fkind[lb * 6 + f]; int& arg = r.H.farg[lb * 6 + f];
        r.H.ftype[lb * 6 + f] = 0;
        if (q < 0 || q >= ext[d]) { kind = FK_PHYS; arg = -1; r.H.ftype[lb * 6 + f] = bc_types[f]; continue; }
        const int ng = g + ((f & 1) ? stride[d] : -stride[d]);
        if (owner(ng) == me) { kind = FK_SAME; arg = local_of[ng]; }
        else { kind = FK_PLANE; arg = r.n_planes; sends[owner(ng)].push_back({local_of[ng], f, me, r.n_planes}); ++r.n_planes; }
      }
      for (int v = 0; v < 4; ++v)
        std::copy(G.var[v].begin() + (size_t)g * G.L.bs, G.var[v].begin() + (size_t)(g + 1) * G.L.bs, r.H.var[v].begin() + (size_t)lb * G.L.bs);
      std::copy(G.eps[0].begin() + (size_t)g * G.L.bs, G.eps[0].begin() + (size_t)(g + 1) * G.L.bs, r.H.eps[0].begin() + (size_t)lb * G.L.bs);
      std::copy(G.epsf[0].begin() + (size_t)g * 6 * nc * nc, G.epsf[0].begin() + (size_t)(g + 1) * 6 * nc * nc, r.H.epsf[0].begin() + (size_t)lb * 6 * nc * nc);
    }
    r.H.fkind_old = r.H.fkind; r.H.farg_old = r.H.farg;
    r.H.L.fkind_old = r.H.fkind_old.data(); r.H.L.farg_old = r.H.farg_old.data();
    r.pool.assign((size_t)2 * std::max(1, r.n_planes) * nc * nc, -7.0);
    r.flags.assign(n_ranks, 0ull);
  }
  g_planes = save_planes;
  for (int me = 0; me < n_ranks; ++me) {
    RankLevel& r = R[me];
    const int nl = (int)r.gid.size();
    LevelDev& L = r.H.L;
    L.gpl = L.gpl_old = r.pool.data(); L.gpl_stride = (long long)std::max(1, r.n_planes) * nc * nc; L.gpl_sel = &r.epoch;
    // partition-face boxes and their items, as sync_meta lays them out
    r.slot_of.assign(nl, -1); r.ptr.assign(1, 0);
    std::vector<int> bnd;
    for (const Send& s : sends[me]) if (r.slot_of[s.box] < 0) { r.slot_of[s.box] = (int)bnd.size(); bnd.push_back(s.box); }
    for (int lb : bnd) {
      for (const Send& s : sends[me]) if (s.box == lb) { r.item.push_back(s.f); r.item.push_back(s.to); r.item.push_back(s.plane); }
      r.ptr.push_back((int)r.item.size() / 3);
    }
    for (int lb : bnd) r.launch.push_back(lb);
    for (int lb = 0; lb < nl; ++lb) if (r.slot_of[lb] < 0) r.launch.push_back(lb);
    PutDev& U = L.put;
    std::memset(&U, 0, sizeof(U));
    U.n_bnd = (int)bnd.size(); U.me = me; U.nranks = n_ranks; U.channel = 0;
    std::vector<int> peers;
    for (const Send& s : sends[me]) if (std::find(peers.begin(), peers.end(), s.to) == peers.end()) peers.push_back(s.to);
    if (peers.size() > 8) return -3;
    U.n_peers = (int)peers.size();
    for (int q = 0; q < U.n_peers; ++q) U.peers[q] = peers[q];
    for (int q = 0; q < n_ranks; ++q) { r.peer_base.push_back(R[q].pool.data()); r.peer_stride.push_back((long long)std::max(1, R[q].n_planes) * nc * nc); r.peer_flags.push_back(R[q].flags.data()); }
    U.slot_of = r.slot_of.data(); U.ptr = r.ptr.data(); U.item = r.item.data();
    U.peer_base = r.peer_base.data(); U.peer_stride = r.peer_stride.data(); U.peer_flags = r.peer_flags.data();
    U.flags = r.flags.data(); U.epoch = &r.epoch; U.done = &r.done; U.err = &r.err;
  }
  // the exchange that precedes the first fused half-sweep (S_HALO): current faces into copy epoch & 1 = 0 of the pools
  for (int me = 0; me < n_ranks; ++me)
    for (const Send& s : sends[me]) {
      const int g = R[me].gid[s.box], layer = (s.f & 1) ? 0 : nc - 1;
      double* o = R[s.to].pool.data() + (size_t)s.plane * nc * nc;
      for (int t2 = 0; t2 < nc; ++t2) for (int t1 = 0; t1 < nc; ++t1) {
        const int i = s.f < 2 ? layer : t1, j = s.f < 2 ? t1 : (s.f < 4 ? layer : t2), k = s.f < 4 ? t2 : layer;
        o[t2 * nc + t1] = G.var[0][(size_t)g * G.L.bs + cidx(G.L, i, j, k)];
      }
    }
  int bad = 0;
  for (int n = 0; n < 4; ++n) {
    for (int b = 0; b < nb; ++b) do_smooth<OP>(G.L, P, b, n & 1, 0, 1);
    for (int me = 0; me < n_ranks; ++me) {
      RankLevel& r = R[me];
      if (!omgf::smooth(r.H.L, P, n & 1, true, bnd_first ? r.launch.data() : nullptr, r.H.L.nb, 1, nullptr)) return -1;
      if (r.epoch != (unsigned long long)(n + 1) && r.H.L.put.n_bnd > 0) ++bad;      // the last partition-face CTA bumps the epoch
      if (r.err) ++bad;
    }
    for (int me = 0; me < n_ranks; ++me)
      for (size_t lb = 0; lb < R[me].gid.size(); ++lb)
        bad += std::memcmp(&R[me].H.var[0][lb * G.L.bs], &G.var[0][(size_t)R[me].gid[lb] * G.L.bs], sizeof(double) * G.L.bs) != 0;
  }
  // correct_children with the in-kernel put of the corrected faces (k_corr_prolong_fast, fused = 2: one fence per CTA), then
  // two more fused half-sweeps, which read the planes that kernel sent.  Box b's parent is coarse box b (an isolated box
  // between physical boundaries, so that a rank's coarse level is a plain copy of its share), its octant varies with b.
  if ((nc == 16 || nc == 32) && (OP == OP_LPL || OP == OP_HELMH)) {
    const int hc = nc / 2;
    auto make_coarse = [&](HostLevel& C, int n, const std::vector<int>* gid) {
      const int save = g_planes; g_planes = 0;
      build_level(C, nc, n, 1, 1, bc_types, 0.02);
      g_planes = save;
      set_fac(C, OP, lambda);
      for (int q = 0; q < n * 6; ++q) { C.fkind[q] = FK_PHYS; C.farg[q] = -1; C.ftype[q] = bc_types[q % 6]; }
      C.fkind_old = C.fkind; C.farg_old = C.farg;
      C.L.fkind_old = C.fkind_old.data(); C.L.farg_old = C.farg_old.data();
      (void)gid;
    };
    HostLevel Gc;
    make_coarse(Gc, nb, nullptr);
    randomise(Gc, seed + 91);
    for (int b = 0; b < nb; ++b) { G.parent[b] = b; G.coff[b] = ((b & 1) * hc) | ((((b >> 1) & 1) * hc) << 10) | ((((b >> 2) & 1) * hc) << 20); }
    std::vector<HostLevel> Rc(n_ranks);
    for (int me = 0; me < n_ranks; ++me) {
      RankLevel& r = R[me];
      const int nl = (int)r.gid.size();
      make_coarse(Rc[me], nl, &r.gid);
      for (int lb = 0; lb < nl; ++lb) {
        const int g = r.gid[lb];
        for (int v = 0; v < 4; ++v)
          std::copy(Gc.var[v].begin() + (size_t)g * Gc.L.bs, Gc.var[v].begin() + (size_t)(g + 1) * Gc.L.bs, Rc[me].var[v].begin() + (size_t)lb * Gc.L.bs);
        r.H.parent[lb] = lb; r.H.coff[lb] = G.coff[g];
      }
    }
    for (int b = 0; b < nb; ++b) do_corr_prolong(G.L, Gc.L, P, b, 0, 1);
    for (int me = 0; me < n_ranks; ++me) {
      RankLevel& r = R[me];
      const unsigned long long e0 = r.epoch;
      if (!omgf::corr_prolong(r.H.L, Rc[me].L, P, true, 2, nullptr)) return -1;
      if (r.epoch != e0 + 1 && r.H.L.put.n_bnd > 0) ++bad;
      if (r.err) ++bad;
    }
    for (int n = 0; n < 2; ++n) {
      for (int b = 0; b < nb; ++b) do_smooth<OP>(G.L, P, b, n & 1, 0, 1);
      for (int me = 0; me < n_ranks; ++me)
        if (!omgf::smooth(R[me].H.L, P, n & 1, true, nullptr, R[me].H.L.nb, 2, nullptr)) return -1;
      for (int me = 0; me < n_ranks; ++me)
        for (size_t lb = 0; lb < R[me].gid.size(); ++lb)
          bad += std::memcmp(&R[me].H.var[0][lb * G.L.bs], &G.var[0][(size_t)R[me].gid[lb] * G.L.bs], sizeof(double) * G.L.bs) != 0;
    }
  }
  return bad;
}


// mg_smoother_gs: ghost snapshot + a plain lexicographic sweep (i fastest, the reference's loop order) against the
// shared-memory staged hyperplane sweep k_gs_sweep_fast (one CTA per box), two sweeps
template <int OP>
int gs_level(int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda, uint64_t seed) {
  HostLevel G, F;
  build_level(G, nc, nbx, nby, nbz, bc_types, 0.01);
  build_level(F, nc, nbx, nby, nbz, bc_types, 0.01);
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  P.smoother = 1;
  set_fac(G, OP, lambda); set_fac(F, OP, lambda);
  randomise(G, seed);
  copy_data(F, G);
  std::vector<double> snap_g((size_t)G.L.nb * 6 * nc * nc, -7.0), snap_f(snap_g);
  G.L.faces[0] = snap_g.data(); F.L.faces[0] = snap_f.data();
  int bad = 0;
  for (int n = 0; n < 2; ++n) {
    for (int b = 0; b < G.L.nb; ++b) do_gs_snapshot(G.L, P, b, 0, 1);
    for (int b = 0; b < G.L.nb; ++b)
      for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i)
        G.L.var[0][(size_t)b * G.L.bs + cidx(G.L, i, j, k)] = gs_at<OP, GM_STORED>(G.L, P, b, i, j, k);
    for (int b = 0; b < F.L.nb; ++b) do_gs_snapshot(F.L, P, b, 0, 1);
    if (!omgf::gs_sweep(F.L, P, nullptr)) return -1;
    bad += differing(G, F) ? 1 : 0;
  }
  return bad;
}

}  // namespace

extern "C" {
// fused half-sweeps with the in-kernel exchange on n_ranks emulated ranks against the single-rank general path
int fe_ranks(int op, int nc, int nbx, int nby, int nbz, int n_ranks, int bnd_first, const int* bc_types, const double* bc_values,
             double lambda, unsigned long long seed) {
  switch (op) {
    case OP_LPL: return ranks_case<OP_LPL>(nc, nbx, nby, nbz, n_ranks, bnd_first, bc_types, bc_values, lambda, seed);
    case OP_VLPL: return ranks_case<OP_VLPL>(nc, nbx, nby, nbz, n_ranks, bnd_first, bc_types, bc_values, lambda, seed);
    case OP_HELMH: return ranks_case<OP_HELMH>(nc, nbx, nby, nbz, n_ranks, bnd_first, bc_types, bc_values, lambda, seed);
    default: return ranks_case<OP_VHELMH>(nc, nbx, nby, nbz, n_ranks, bnd_first, bc_types, bc_values, lambda, seed);
  }
}
int fe_gs_level(int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda, unsigned long long seed) {
  return op == OP_HELMH ? gs_level<OP_HELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed)
                        : gs_level<OP_LPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
}
// mismatching comparisons between the fast kernels and the general path (0 = same bits); -1: a kernel did not apply
int fe_one_level(int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda,
                 unsigned long long seed) {
  switch (op) {
    case OP_LPL: return one_level<OP_LPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_VLPL: return one_level<OP_VLPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_HELMH: return one_level<OP_HELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_AHELMH: return one_level<OP_AHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    default: return one_level<OP_VHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
  }
}
int fe_two_levels(int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda,
                  unsigned long long seed) {
  switch (op) {
    case OP_LPL: return two_levels<OP_LPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_VLPL: return two_levels<OP_VLPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_HELMH: return two_levels<OP_HELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_AHELMH: return two_levels<OP_AHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    default: return two_levels<OP_VHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
  }
}
// 0 / 1 / 2: never / on big levels / always use the persistent ring smoother (omgf::set_ring)
void fe_set_ring(int mode) { omgf::set_ring(mode); }
// checks of the emulated TMA / mbarrier / shared-memory rules that failed so far
int fe_violations() { return emu::g_violations.load(); }
}

// ---- the single-CTA tail (k_tail, omg_tail.cuh) on a hierarchy of coarse levels ---------------------------------------
// n_levels levels of nbx x nby x nbz boxes of 2^3, 4^3, 8^3, 16^3 cells (the single-box / few-box coarsening levels below
// first_normal_lvl: box b of a level is refined into box b of the next, src/m_build_tree.f90:60-110), a V(2,2) op list as
// the program builder emits it (omg_core.cu Builder::vcycle) incl. the coarse solve, run by 1024 (as launched) or fewer emulated threads with and
// without shared-memory staging, against the same ops applied one after the other by the general per-cell functions.
namespace {

struct Hier {
  std::vector<HostLevel> H;
  std::vector<std::vector<int>> plist;
  std::vector<LevelDev> dev;
};
void build_hier(Hier& Q, int n_levels, int nbx, int nby, int nbz, const int* bc_types, int op, double lambda, int simple) {
  Q.H.resize(n_levels); Q.plist.resize(n_levels);
  for (int l = 0; l < n_levels; ++l) {
    HostLevel& H = Q.H[l];
    build_level(H, 2 << l, nbx, nby, nbz, bc_types, 0.08 / (1 << l));
    set_fac(H, op, lambda);
    H.L.lvl = l; H.L.simple = simple;
    for (int b = 0; b < H.L.nb; ++b) { H.parent[b] = l > 0 ? b : -1; H.coff[b] = 0; }
    if (l + 1 < n_levels) { for (int b = 0; b < H.L.nb; ++b) Q.plist[l].push_back(b); H.L.parents_list = Q.plist[l].data(); H.L.n_parents = H.L.nb; }
  }
}
void sync_dev(Hier& Q) { Q.dev.clear(); for (HostLevel& H : Q.H) Q.dev.push_back(H.L); }

std::vector<TailOp> vcycle_ops(int n_levels, int coarse_sub, int max_coarse) {
  std::vector<TailOp> ops;
  const int top = n_levels - 1;
  for (int l = top; l >= 1; --l) {
    for (int n = 0; n < 4; ++n) ops.push_back({TOP_SMOOTH, l, n & 1, 0});
    ops.push_back({TOP_RESID, l, -1, 0});
    ops.push_back({TOP_RESTRICT, l, 0, 3});
    ops.push_back({TOP_CRHS, l - 1, 0, 0});
  }
  ops.push_back({TOP_COARSE_SOLVE, 0, coarse_sub, max_coarse});
  for (int l = 1; l <= top; ++l) {
    ops.push_back({TOP_CORR_PROLONG, l, 0, 0});
    for (int n = 0; n < 4; ++n) ops.push_back({TOP_SMOOTH, l, n & 1, 0});
  }
  for (int l = 0; l <= top; ++l) ops.push_back({TOP_RESID, l, SCAL_MAXRES, 0});
  return ops;
}

template <int OP>
void reference_ops(Hier& Q, const ParamsDev& P, const std::vector<TailOp>& ops, double* scal) {
  auto resid_max = [&](const LevelDev& L) { double m = 0.0; for (int b = 0; b < L.nb; ++b) m = fmax(m, do_resid<OP>(L, P, b, 0, 1)); return m; };
  for (const TailOp& op : ops) {
    const LevelDev& L = Q.H[op.l].L;
    switch (op.type) {
      case TOP_SMOOTH: for (int b = 0; b < L.nb; ++b) do_smooth<OP>(L, P, b, op.a, 0, 1); break;
      case TOP_RESID: { const double m = resid_max(L); if (op.a >= 0 && m > scal[op.a]) scal[op.a] = m; } break;
      case TOP_RESTRICT: for (int b = 0; b < L.nb; ++b) { do_restrict(L, Q.H[op.l - 1].L, op.a, b, 0, 1); if (op.b >= 0) do_restrict(L, Q.H[op.l - 1].L, op.b, b, 0, 1); } break;
      case TOP_CRHS: for (int q = 0; q < L.n_parents; ++q) do_crhs<OP>(L, P, L.parents_list[q], 0, 1); break;
      case TOP_CORR_PROLONG: for (int b = 0; b < L.nb; ++b) do_corr_prolong(L, Q.H[op.l - 1].L, P, b, 0, 1); break;
      case TOP_COARSE_SOLVE: {                     // src/m_multigrid.f90:202-208
        const double init_res = resid_max(L);
        int it = 0;
        for (int i = 1; i <= op.b; ++i) {
          for (int n = 1; n <= op.a; ++n) for (int b = 0; b < L.nb; ++b) do_smooth<OP>(L, P, b, (n - 1) & 1, 0, 1);
          const double res = resid_max(L);
          it = i;
          if (res < scal[SCAL_COARSE_REL] * init_res || res < scal[SCAL_COARSE_ABS]) break;
        }
        scal[SCAL_COARSE_ITERS] = (double)it;
      } break;
      default: break;
    }
  }
}

template <int OP>
int tail_case(int n_threads, int n_levels, int nbx, int nby, int nbz, int simple, int staged, const int* bc_types, const double* bc_values, double lambda, uint64_t seed) {
  Hier G, T;
  build_hier(G, n_levels, nbx, nby, nbz, bc_types, OP, lambda, 0);
  build_hier(T, n_levels, nbx, nby, nbz, bc_types, OP, lambda, simple);
  for (int l = 0; l < n_levels; ++l) { randomise(G.H[l], seed + 5 * l); copy_data(T.H[l], G.H[l]); }
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  P.warp_coarse = 1;     // as the product: a single 2^3 box between physical boundaries is solved by warp_coarse_solve (simple levels only)
  const std::vector<TailOp> ops = vcycle_ops(n_levels, 4, 8);
  std::vector<double> sg(SCAL_COUNT, 0.0), st(SCAL_COUNT, 0.0);
  sg[SCAL_COARSE_REL] = st[SCAL_COARSE_REL] = 0.05;   // ends after a few iterations where the coarse problem converges, at the limit where it is singular
  sg[SCAL_COARSE_ABS] = st[SCAL_COARSE_ABS] = 1e-300;
  reference_ops<OP>(G, P, ops, sg.data());
  sync_dev(T);
  size_t stage_bytes = 0;
  if (staged) for (const LevelDev& L : T.dev) stage_bytes += (size_t)4 * L.nb * L.bs * 8 + (size_t)32 * L.nb * sizeof(int);
  if (stage_bytes > (size_t)200 * 1024) return -2;
  const LevelDev* levels = T.dev.data(); const TailOp* po = ops.data(); double* scal = st.data();
  const int n_ops = (int)ops.size(), n_stage = staged ? n_levels : 0;
  emu::launch(1, n_threads, stage_bytes, [&]() { k_tail(levels, P, po, n_ops, scal, n_stage); });
  int bad = 0;
  for (int l = 0; l < n_levels; ++l) bad += differing(G.H[l], T.H[l]);
  bad += std::memcmp(&sg[SCAL_COARSE_ITERS], &st[SCAL_COARSE_ITERS], 8) != 0;
  bad += std::memcmp(&sg[SCAL_MAXRES], &st[SCAL_MAXRES], 8) != 0;
  if (sg[SCAL_COARSE_ITERS] < 1.0) bad += 100;
  return bad;
}

}  // namespace

extern "C" int fe_tail(int op, int n_threads, int n_levels, int nbx, int nby, int nbz, int simple, int staged, const int* bc_types, const double* bc_values, double lambda,
                       unsigned long long seed) {
  switch (op) {
    case OP_LPL: return tail_case<OP_LPL>(n_threads, n_levels, nbx, nby, nbz, simple, staged, bc_types, bc_values, lambda, seed);
    case OP_VLPL: return tail_case<OP_VLPL>(n_threads, n_levels, nbx, nby, nbz, simple, staged, bc_types, bc_values, lambda, seed);
    case OP_HELMH: return tail_case<OP_HELMH>(n_threads, n_levels, nbx, nby, nbz, simple, staged, bc_types, bc_values, lambda, seed);
    case OP_VHELMH: return tail_case<OP_VHELMH>(n_threads, n_levels, nbx, nby, nbz, simple, staged, bc_types, bc_values, lambda, seed);
    default: return tail_case<OP_AHELMH>(n_threads, n_levels, nbx, nby, nbz, simple, staged, bc_types, bc_values, lambda, seed);
  }
}
