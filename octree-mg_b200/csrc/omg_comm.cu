// omg_comm.cu -- transport back ends (see omg_comm.h): NCCL send/recv groups over NVLink, and an
// in-process emulation of several ranks for single-GPU parity tests.
#include "omg_comm.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <nccl.h>   // types only: the library is loaded at run time (see nccl_api)

namespace {
[[noreturn]] void cfail(const std::string& m) { throw std::runtime_error(m); }

// NCCL is bound with dlopen when the first communicator is made, not at link time: a process that also uses
// PyTorch must end up with ONE libnccl.so.2 (torch ships its own, newer than the system one, and fails to
// import if an older copy is already mapped).  OMG_NCCL_LIB names the file; default: the soname, which
// resolves to a copy that is already loaded, else to the system library.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
const NcclApi& nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  static std::string err;
  std::call_once(once, [] {
    const char* path = getenv("OMG_NCCL_LIB");
    void* lib = dlopen(path && *path ? path : "libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!lib) { err = std::string("cannot load NCCL: ") + dlerror(); return; }
    auto sym = [&](const char* n) { void* p = dlsym(lib, n); if (!p && err.empty()) err = std::string("NCCL symbol missing: ") + n; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
  });
  if (!err.empty()) cfail(err);
  return api;
}
#define CCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) cfail(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
#define NCK(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) cfail(std::string(#call) + ": " + nccl_api().GetErrorString(r_)); } while (0)

__global__ void k_reduce_slots(const double* __restrict__ slots, int n, int op, double* out) {
  // fixed rank order: every rank computes the same bits
  double v = slots[0];
  for (int r = 1; r < n; ++r) v = (op == OMG_RED_MAX) ? fmax(v, slots[r]) : v + slots[r];
  *out = v;
}

// ---- peer-mapped puts: signalling ---------------------------------------------------------------------------------
struct PeerList { int n; int peer[31]; };
constexpr int kMaxRanks = 32;
// flags[channel][r]: number of put phases rank r has completed towards me on that channel (written by r through its
// mapping of my flag array).  *epoch: number of put phases I have completed on the channel; all ranks run the same
// sequence of phases, so "flags[p] >= my epoch" means p's puts of the current phase have landed.
// The wait is an acquire (ld.acquire.sys): the peer's puts, released by its st.release.sys of the flag, are visible to
// everything this thread does afterwards; callers follow it with a CTA barrier for the other threads.
// timeout_ns = 0: wait for ever (MPI_Waitall semantics); else a peer that has not arrived after that long raises the
// error word (omg_p2p_err_code: peer + channel) -- the host finds it at its next synchronisation point and poisons the
// handle -- and the waiter returns false, so that the caller leaves its epoch alone.
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
    __nanosleep(100);
  }
}
__device__ __forceinline__ void signal_flag(unsigned long long* f, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(v) : "memory");
}
// signal = 1: raise my flag at the peers (my puts precede this kernel in stream order), wait for theirs, bump the epoch.
// signal = 0: only wait for the peers' flags of the current epoch (puts that were signalled from inside a kernel).
__global__ void k_signal_wait(PeerList pl, int channel, int me, int nranks, unsigned long long* const* __restrict__ peer_flags,
                              unsigned long long* flags, unsigned long long* epoch, int* err, int signal, unsigned long long timeout_ns) {
  const int t = threadIdx.x;
  const unsigned long long e = *epoch + (signal ? 1ull : 0ull);
  __shared__ int ok;
  if (t == 0) ok = 1;
  __syncthreads();
  if (t < pl.n) {
    const int p = pl.peer[t];
    if (signal) signal_flag(peer_flags[p] + channel * nranks + me, e);     // release: my puts precede this kernel in stream order
    if (!wait_flag(flags + channel * nranks + p, e, err, timeout_ns, omg_p2p_err_code(p, channel))) ok = 0;
  }
  __syncthreads();
  if (t == 0 && signal && ok) *epoch = e;       // a failed wait leaves the epoch alone: the handle is poisoned on the host
}
__global__ void k_bump_epoch(unsigned long long* epoch) { *epoch = *epoch + 1; }
__global__ void k_add_vec(double* __restrict__ acc, const double* __restrict__ x, size_t n, int first) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc[i] = first ? x[i] : acc[i] + x[i];
}
}  // namespace

struct omg_local_group {
  int n = 0, refs = 0;
  std::mutex m;
  std::condition_variable cv;
  int count = 0;
  unsigned long gen = 0;
  bool broken = false;
  std::vector<omg_comm*> comms;
  std::vector<void*> share_tbl;             // omg_comm_share: the pointer every emulated rank contributes

  // all emulated ranks meet here; a rank that failed (or never arrives) breaks the group instead of hanging it
  void barrier() {
    std::unique_lock<std::mutex> lk(m);
    if (broken) cfail("local rank group is broken (another emulated rank failed)");
    const unsigned long my = gen;
    if (++count == n) { count = 0; ++gen; cv.notify_all(); return; }
    if (!cv.wait_for(lk, std::chrono::seconds(120), [&] { return gen != my || broken; })) {
      broken = true; cv.notify_all();
      cfail("local rank group: barrier timed out (ranks must call the same operations)");
    }
    if (broken && gen == my) cfail("local rank group is broken (another emulated rank failed)");
  }
  void abort_group() { std::lock_guard<std::mutex> lk(m); broken = true; cv.notify_all(); }
};

struct omg_comm {
  int rank = 0, n = 1, device = 0;
  ncclComm_t nccl = nullptr;
  omg_local_group* g = nullptr;
  // local back end
  std::vector<OmgXfer> pub;                 // segments of the exchange in flight, visible to the peers
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  double* d_stage = nullptr;                // my contribution to an allreduce
  double* d_gather = nullptr;               // [n] contributions of all ranks
  double* d_vec = nullptr; size_t n_vec = 0;   // my contribution to a vector all-reduce
  // peer-mapped puts (NCCL back end)
  bool p2p = false;
  unsigned long long* d_flags = nullptr;    // [OMG_P2P_CHANNELS][n], written by the peers
  unsigned long long** d_peer_flags = nullptr;
  int* d_err = nullptr;                     // bounded-wait error word: mapped pinned host memory (h_err is the host view),
  volatile int* h_err = nullptr;            // so the host reads it without touching the device
  unsigned long long timeout_ns = 0;
  std::vector<void*> ipc_opened;
  size_t n_flag_maps = 0;                   // the first entries of ipc_opened: the peers' flag arrays (kept by omg_comm_unshare)
  void* ipc_scratch = nullptr;
};

// error word of the bounded waits + the bound itself (OMG_P2P_TIMEOUT_S seconds, default 120, 0 = wait for ever)
static void alloc_err(omg_comm* c) {
  int* hp = nullptr;
  CCK(cudaHostAlloc(&hp, sizeof(int), cudaHostAllocMapped));
  *hp = 0;
  c->h_err = hp;
  CCK(cudaHostGetDevicePointer(&c->d_err, hp, 0));
  const char* e = getenv("OMG_P2P_TIMEOUT_S");
  const double sec = (e && *e) ? atof(e) : 120.0;
  c->timeout_ns = sec > 0 ? (unsigned long long)(sec * 1e9) : 0ull;
}

int omg_comm_unique_id(char out[128]) {
  static_assert(sizeof(ncclUniqueId) <= 128, "ncclUniqueId larger than the ABI buffer");
  ncclUniqueId id;
  try { if (nccl_api().GetUniqueId(&id) != ncclSuccess) return 1; } catch (...) { return 2; }
  std::memset(out, 0, 128);
  std::memcpy(out, &id, sizeof(id));
  return 0;
}

omg_comm* omg_comm_create_nccl(const char uid[128], int rank, int nranks, int device) {
  CCK(cudaSetDevice(device));
  ncclUniqueId id;
  std::memcpy(&id, uid, sizeof(id));
  omg_comm* c = new omg_comm();
  c->rank = rank; c->n = nranks; c->device = device;
  ncclResult_t r;
  try { r = nccl_api().CommInitRank(&c->nccl, nranks, id, rank); } catch (...) { delete c; throw; }
  if (r != ncclSuccess) { delete c; cfail(std::string("ncclCommInitRank: ") + nccl_api().GetErrorString(r)); }
  // flag array for peer-mapped puts, shared with every rank; OMG_P2P=0 keeps everything on ncclSend/ncclRecv
  const char* env = getenv("OMG_P2P");
  if (nranks <= kMaxRanks && !(env && *env == '0')) {
    try {
      const size_t nf = (size_t)OMG_P2P_CHANNELS * nranks;
      CCK(cudaMalloc(&c->d_flags, nf * sizeof(unsigned long long)));
      CCK(cudaMemset(c->d_flags, 0, nf * sizeof(unsigned long long)));
      alloc_err(c);
      CCK(cudaMalloc(&c->d_peer_flags, nranks * sizeof(void*)));
      std::vector<void*> pf(nranks, nullptr);
      c->p2p = true;                                   // omg_comm_share clears it again (on every rank) if IPC is unavailable
      omg_comm_share(c, c->d_flags, pf.data());
      if (c->p2p) CCK(cudaMemcpy(c->d_peer_flags, pf.data(), nranks * sizeof(void*), cudaMemcpyHostToDevice));
      c->n_flag_maps = c->ipc_opened.size();
    } catch (...) { c->p2p = false; }
  }
  return c;
}

omg_local_group* omg_local_group_create(int nranks) {
  omg_local_group* g = new omg_local_group();
  g->n = nranks; g->comms.assign(nranks, nullptr); g->share_tbl.assign(nranks, nullptr);
  return g;
}

omg_comm* omg_comm_create_local(omg_local_group* g, int rank, int device) {
  CCK(cudaSetDevice(device));
  omg_comm* c = new omg_comm();
  c->rank = rank; c->n = g->n; c->device = device; c->g = g;
  CCK(cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming));
  CCK(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
  CCK(cudaMalloc(&c->d_stage, sizeof(double)));
  CCK(cudaMalloc(&c->d_gather, sizeof(double) * g->n));
  std::lock_guard<std::mutex> lk(g->m);
  g->comms[rank] = c; g->refs++;
  return c;
}

// Ordered release of the peer mappings (default; OMG_ORDERED_TEARDOWN=0 switches it off).  Freeing memory a peer still
// has mapped is undefined (CUDA IPC rules) and the ranks of a job leave at different moments, so when every rank
// deallocates (the reference's mg_deallocate_storage is called by all of them) the order is: all ranks arrive (bounded
// flag wait over the mappings themselves -- a missing rank makes the others give up after 20 s instead of hanging) ->
// everyone closes what it imported -> NCCL barrier -> the exported memory is freed by the callers.  Returns true when
// the ranks met.
static bool teardown_rendezvous(omg_comm* c) {
  const char* e = getenv("OMG_ORDERED_TEARDOWN");
  if (e && *e && atoi(e) == 0) return false;
  if (!c->nccl || !c->p2p || c->n < 2 || c->n > kMaxRanks || !c->d_peer_flags || !c->d_flags || !c->d_err) return false;
  if (*c->h_err) return false;   // an exchange already failed
  unsigned long long* ep = nullptr;
  if (cudaMalloc(&ep, sizeof(unsigned long long)) != cudaSuccess) return false;
  cudaMemset(ep, 0, sizeof(unsigned long long));
  PeerList pl; pl.n = 0;
  for (int r = 0; r < c->n; ++r) if (r != c->rank) pl.peer[pl.n++] = r;
  const unsigned long long bound = 20000000000ull;          // 20 s: ranks that deallocate together arrive together
  k_signal_wait<<<1, 32>>>(pl, OMG_P2P_CHANNELS - 1, c->rank, c->n, c->d_peer_flags, c->d_flags, ep, c->d_err, 1, bound);
  bool ok = cudaDeviceSynchronize() == cudaSuccess;
  ok = ok && *c->h_err == 0;
  cudaFree(ep);
  return ok;
}

void omg_comm_destroy(omg_comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  const char* dbg_e = getenv("OMG_DEBUG_TEARDOWN");
  const bool dbg = dbg_e && *dbg_e && atoi(dbg_e) != 0;
  auto say = [&](const char* what) { if (dbg) { std::fprintf(stderr, "[omg teardown rank %d] %s\n", c->rank, what); std::fflush(stderr); } };
  say("rendezvous");
  const bool met = teardown_rendezvous(c);
  say(met ? "met; closing peer mappings" : "no rendezvous; closing peer mappings");
  for (void* p : c->ipc_opened) cudaIpcCloseMemHandle(p);
  say("closed");
  if (met) {                                 // nobody frees before everybody has closed
    double* d = nullptr;
    if (cudaMalloc(&d, sizeof(double)) == cudaSuccess) {
      cudaMemset(d, 0, sizeof(double));
      if (nccl_api().AllReduce(d, d, 1, ncclDouble, ncclSum, c->nccl, 0) == ncclSuccess) cudaDeviceSynchronize();
      cudaFree(d);
    }
  }
  if (c->d_flags) cudaFree(c->d_flags);
  if (c->ipc_scratch) cudaFree(c->ipc_scratch);
  if (c->h_err) cudaFreeHost((void*)c->h_err);
  if (c->d_peer_flags) cudaFree(c->d_peer_flags);
  say("freed; destroying the NCCL communicator");
  if (c->nccl) nccl_api().CommDestroy(c->nccl);
  say("done");
  if (c->g) {
    c->g->abort_group();                    // a group cannot continue without one of its ranks
    if (c->ev_ready) cudaEventDestroy(c->ev_ready);
    if (c->ev_done) cudaEventDestroy(c->ev_done);
    if (c->d_stage) cudaFree(c->d_stage);
    if (c->d_gather) cudaFree(c->d_gather);
    if (c->d_vec) cudaFree(c->d_vec);
    bool last;
    { std::lock_guard<std::mutex> lk(c->g->m); c->g->comms[c->rank] = nullptr; last = (--c->g->refs == 0); }
    if (last) delete c->g;
  }
  delete c;
}

int omg_comm_rank(const omg_comm* c) { return c->rank; }
int omg_comm_size(const omg_comm* c) { return c->n; }
bool omg_comm_graph_safe(const omg_comm* c) { return c->nccl != nullptr; }

void omg_comm_exchange(omg_comm* c, const OmgXfer* segs, int nsegs, cudaStream_t s) {
  if (c->nccl) {
    // sort_and_transfer_buffers: Isend + Irecv per peer, then Waitall (src/m_communication.f90:48-64);
    // here one fused group on the stream, completion is stream order
    const NcclApi& N = nccl_api();
    if (nsegs == 0) return;
    NCK(N.GroupStart());
    for (int i = 0; i < nsegs; ++i) {
      if (segs[i].send_n) NCK(N.Send(segs[i].send, segs[i].send_n, ncclDouble, segs[i].peer, c->nccl, s));
      if (segs[i].recv_n) NCK(N.Recv(segs[i].recv, segs[i].recv_n, ncclDouble, segs[i].peer, c->nccl, s));
    }
    NCK(N.GroupEnd());
    return;
  }
  omg_local_group* g = c->g;
  try {
    c->pub.assign(segs, segs + nsegs);
    CCK(cudaEventRecord(c->ev_ready, s));                    // my send buffers are packed once this fires
    g->barrier();
    for (int i = 0; i < nsegs; ++i) {
      if (!segs[i].recv_n) continue;
      omg_comm* p = g->comms[segs[i].peer];
      if (!p) cfail("local exchange: peer rank is gone");
      const OmgXfer* src = nullptr;
      for (const OmgXfer& x : p->pub) if (x.peer == c->rank) src = &x;
      if (!src || src->send_n != segs[i].recv_n) cfail("local exchange: send/recv plans of two ranks do not match");
      CCK(cudaStreamWaitEvent(s, p->ev_ready, 0));
      CCK(cudaMemcpyAsync(segs[i].recv, src->send, segs[i].recv_n * sizeof(double), cudaMemcpyDefault, s));
    }
    CCK(cudaEventRecord(c->ev_done, s));                     // I have read my peers' send buffers
    g->barrier();
    for (int i = 0; i < nsegs; ++i)
      if (segs[i].send_n && g->comms[segs[i].peer]) CCK(cudaStreamWaitEvent(s, g->comms[segs[i].peer]->ev_done, 0));
    g->barrier();                                            // events may be re-recorded only after everyone waited on them
  } catch (...) { g->abort_group(); throw; }
}

void omg_comm_allreduce(omg_comm* c, double* d_val, int op, cudaStream_t s) {
  if (c->nccl) {
    NCK(nccl_api().AllReduce(d_val, d_val, 1, ncclDouble, op == OMG_RED_MAX ? ncclMax : ncclSum, c->nccl, s));
    return;
  }
  omg_local_group* g = c->g;
  try {
    CCK(cudaMemcpyAsync(c->d_stage, d_val, sizeof(double), cudaMemcpyDeviceToDevice, s));
    CCK(cudaEventRecord(c->ev_ready, s));
    g->barrier();
    for (int r = 0; r < c->n; ++r) {
      omg_comm* p = g->comms[r];
      if (!p) cfail("local allreduce: peer rank is gone");
      CCK(cudaStreamWaitEvent(s, p->ev_ready, 0));
      CCK(cudaMemcpyAsync(c->d_gather + r, p->d_stage, sizeof(double), cudaMemcpyDefault, s));
    }
    k_reduce_slots<<<1, 1, 0, s>>>(c->d_gather, c->n, op, d_val);
    CCK(cudaEventRecord(c->ev_done, s));
    g->barrier();
    for (int r = 0; r < c->n; ++r) if (g->comms[r]) CCK(cudaStreamWaitEvent(s, g->comms[r]->ev_done, 0));
    g->barrier();
  } catch (...) { g->abort_group(); throw; }
}

void omg_comm_allreduce_sum_vec(omg_comm* c, double* d_val, size_t n, cudaStream_t s) {
  if (n == 0) return;
  if (c->nccl) { NCK(nccl_api().AllReduce(d_val, d_val, n, ncclDouble, ncclSum, c->nccl, s)); return; }
  omg_local_group* g = c->g;
  try {
    if (c->n_vec < n) { if (c->d_vec) cudaFree(c->d_vec); CCK(cudaMalloc(&c->d_vec, n * sizeof(double))); c->n_vec = n; }
    CCK(cudaMemcpyAsync(c->d_vec, d_val, n * sizeof(double), cudaMemcpyDeviceToDevice, s));
    CCK(cudaEventRecord(c->ev_ready, s));
    g->barrier();
    const int grid = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
    for (int r = 0; r < c->n; ++r) {
      omg_comm* p = g->comms[r];
      if (!p || p->n_vec < n) cfail("local allreduce: peer rank is gone or disagrees on the size");
      CCK(cudaStreamWaitEvent(s, p->ev_ready, 0));
      k_add_vec<<<grid, 256, 0, s>>>(d_val, p->d_vec, n, r == 0);
    }
    CCK(cudaEventRecord(c->ev_done, s));
    g->barrier();
    for (int r = 0; r < c->n; ++r) if (g->comms[r]) CCK(cudaStreamWaitEvent(s, g->comms[r]->ev_done, 0));
    g->barrier();
  } catch (...) { g->abort_group(); throw; }
}

// ---- peer-mapped puts ------------------------------------------------------------------------------------------------
bool omg_comm_p2p(const omg_comm* c) { return c->g ? true : c->p2p; }

void omg_comm_share(omg_comm* c, void* ptr, void** peer_ptrs) {
  if (c->g) {
    omg_local_group* g = c->g;
    try {
      { std::lock_guard<std::mutex> lk(g->m); g->share_tbl[c->rank] = ptr; }
      g->barrier();
      for (int r = 0; r < c->n; ++r) {
        peer_ptrs[r] = g->share_tbl[r];
        if (r != c->rank && g->comms[r] && g->comms[r]->device != c->device) {
          cudaError_t e = cudaDeviceEnablePeerAccess(g->comms[r]->device, 0);
          if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) cfail("local share: no peer access between the devices of two emulated ranks");
          cudaGetLastError();
        }
      }
      g->barrier();
    } catch (...) { g->abort_group(); throw; }
    return;
  }
  // NCCL back end: CUDA IPC handles through ncclAllGather.  Every rank must reach the same verdict, so the handle
  // travels with a validity byte and the mapping result is agreed on with a second gather.
  const NcclApi& N = nccl_api();
  if (!c->p2p) { for (int r = 0; r < c->n; ++r) peer_ptrs[r] = (r == c->rank) ? ptr : nullptr; return; }
  constexpr size_t kRec = 128;
  static_assert(sizeof(cudaIpcMemHandle_t) < kRec, "IPC handle larger than the record");
  std::vector<unsigned char> rec((size_t)c->n * kRec, 0);
  cudaIpcMemHandle_t hd;
  const bool got = cudaIpcGetMemHandle(&hd, ptr) == cudaSuccess;
  if (!got) cudaGetLastError();
  std::memcpy(&rec[(size_t)c->rank * kRec], &hd, sizeof(hd));
  rec[(size_t)c->rank * kRec + kRec - 1] = got ? 1 : 0;
  unsigned char* d = nullptr;
  CCK(cudaMalloc(&d, rec.size()));
  auto gather = [&]() {
    CCK(cudaMemcpy(d + (size_t)c->rank * kRec, &rec[(size_t)c->rank * kRec], kRec, cudaMemcpyHostToDevice));
    NCK(N.AllGather(d + (size_t)c->rank * kRec, d, kRec, ncclChar, c->nccl, 0));
    CCK(cudaStreamSynchronize(0));
    CCK(cudaMemcpy(rec.data(), d, rec.size(), cudaMemcpyDeviceToHost));
  };
  gather();
  bool ok = true;
  for (int r = 0; r < c->n; ++r) ok = ok && rec[(size_t)r * kRec + kRec - 1] == 1;
  std::vector<void*> opened;
  if (ok) {
    for (int r = 0; r < c->n && ok; ++r) {
      if (r == c->rank) { peer_ptrs[r] = ptr; continue; }
      cudaIpcMemHandle_t hr;
      std::memcpy(&hr, &rec[(size_t)r * kRec], sizeof(hr));
      void* q = nullptr;
      if (cudaIpcOpenMemHandle(&q, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
      peer_ptrs[r] = q; opened.push_back(q);
    }
  }
  rec[(size_t)c->rank * kRec + kRec - 1] = ok ? 1 : 0;      // agree: did every rank map every peer?
  gather();
  for (int r = 0; r < c->n; ++r) ok = ok && rec[(size_t)r * kRec + kRec - 1] == 1;
  cudaFree(d);
  if (!ok) {
    for (void* q : opened) cudaIpcCloseMemHandle(q);
    c->p2p = false;
    for (int r = 0; r < c->n; ++r) peer_ptrs[r] = (r == c->rank) ? ptr : nullptr;
    return;
  }
  c->ipc_opened.insert(c->ipc_opened.end(), opened.begin(), opened.end());
}

void omg_comm_unshare(omg_comm* c) {
  if (c->g) { c->g->barrier(); return; }       // emulated ranks share plain pointers: nothing to close, only the meeting point
  if (!c->nccl) return;
  CCK(cudaDeviceSynchronize());
  for (size_t i = c->n_flag_maps; i < c->ipc_opened.size(); ++i) cudaIpcCloseMemHandle(c->ipc_opened[i]);
  c->ipc_opened.resize(std::min(c->n_flag_maps, c->ipc_opened.size()));
  double* d = nullptr;
  CCK(cudaMalloc(&d, sizeof(double)));
  CCK(cudaMemset(d, 0, sizeof(double)));
  NCK(nccl_api().AllReduce(d, d, 1, ncclDouble, ncclSum, c->nccl, 0));
  CCK(cudaDeviceSynchronize());
  cudaFree(d);
}

void omg_comm_signal_wait(omg_comm* c, int channel, const int* peers, int npeers, unsigned long long* epoch, int signal, cudaStream_t s) {
  if (channel < 0 || channel >= OMG_P2P_CHANNELS) cfail("signal_wait: channel out of range");
  if (c->g) {
    omg_local_group* g = c->g;
    try {
      CCK(cudaEventRecord(c->ev_ready, s));                  // my puts are complete once this fires
      g->barrier();
      for (int i = 0; i < npeers; ++i) {
        omg_comm* p = g->comms[peers[i]];
        if (!p) cfail("local signal_wait: peer rank is gone");
        CCK(cudaStreamWaitEvent(s, p->ev_ready, 0));
      }
      if (signal) k_bump_epoch<<<1, 1, 0, s>>>(epoch);
      g->barrier();                                          // events may be re-recorded only after everyone waited on them
    } catch (...) { g->abort_group(); throw; }
    return;
  }
  if (!c->p2p) cfail("signal_wait without peer-mapped memory");
  if (npeers > 31) cfail("signal_wait: too many peers");
  PeerList pl; pl.n = npeers;
  for (int i = 0; i < npeers; ++i) pl.peer[i] = peers[i];
  k_signal_wait<<<1, 32, 0, s>>>(pl, channel, c->rank, c->n, c->d_peer_flags, c->d_flags, epoch, c->d_err, signal, c->timeout_ns);
  CCK(cudaGetLastError());
}

void omg_comm_p2p_ptrs(omg_comm* c, unsigned long long* const** peer_flags, unsigned long long** flags, int** err) {
  if (c->g) {
    // emulated ranks: in-kernel signalling writes to scratch slots nobody waits on (ordering comes from the host barrier)
    if (!c->d_flags) {
      const size_t nf = (size_t)OMG_P2P_CHANNELS * c->n;
      CCK(cudaMalloc(&c->d_flags, nf * sizeof(unsigned long long)));
      CCK(cudaMemset(c->d_flags, 0xff, nf * sizeof(unsigned long long)));    // "already arrived": waits return at once
      alloc_err(c);
      CCK(cudaMalloc(&c->d_peer_flags, c->n * sizeof(void*)));
      unsigned long long* scratch = nullptr;
      CCK(cudaMalloc(&scratch, nf * sizeof(unsigned long long)));
      c->ipc_scratch = scratch;
      std::vector<unsigned long long*> pf(c->n, scratch);
      CCK(cudaMemcpy(c->d_peer_flags, pf.data(), c->n * sizeof(void*), cudaMemcpyHostToDevice));
    }
  }
  *peer_flags = c->d_peer_flags; *flags = c->d_flags; *err = c->d_err;
}

int omg_comm_p2p_error(omg_comm* c) { return c->h_err ? *c->h_err : 0; }
unsigned long long omg_comm_p2p_timeout_ns(const omg_comm* c) { return c->timeout_ns; }
