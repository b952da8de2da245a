// omg_comm.h -- inter-rank transport (replaces sort_and_transfer_buffers, src/m_communication.f90:37-66,
// and the two scalar MPI_Allreduce calls of src/m_multigrid.f90:232-233,255-256).
//
// Two back ends behind one interface:
//  * NCCL: one process per GPU; one ncclSend/ncclRecv pair per peer inside a group, on the caller's stream.
//  * local: several emulated ranks inside one process (one host thread per rank, any GPU assignment);
//    the exchange is a device-to-device copy from the peer's send buffer, ordered with CUDA events and a
//    host barrier.  Used by the parity tests on a single GPU; no kernel ever waits on another kernel.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

struct omg_comm;

// one peer's part of an exchange: `send_n` doubles leave from `send`, `recv_n` doubles arrive in `recv`
struct OmgXfer { int peer; const double* send; size_t send_n; double* recv; size_t recv_n; };

enum { OMG_RED_MAX = 0, OMG_RED_SUM = 1 };

int omg_comm_unique_id(char out[128]);
omg_comm* omg_comm_create_nccl(const char uid[128], int rank, int nranks, int device);   // throws std::runtime_error
// shared state of a group of emulated ranks; each rank then takes its own omg_comm from it
struct omg_local_group;
omg_local_group* omg_local_group_create(int nranks);
omg_comm* omg_comm_create_local(omg_local_group* g, int rank, int device);
void omg_comm_destroy(omg_comm* c);

int omg_comm_rank(const omg_comm* c);
int omg_comm_size(const omg_comm* c);
bool omg_comm_graph_safe(const omg_comm* c);   // may the exchange be captured into a CUDA graph?
// ---- peer-mapped puts (the halo exchange of the smoother) ---------------------------------------------------------
// Instead of send/recv pairs, a rank's pack kernel stores faces straight into the receiver's plane pool through a
// peer mapping of that allocation (NVLink P2P stores); omg_comm_signal_wait then tells the peers "my puts are done" and
// waits for theirs.  NCCL back end: CUDA IPC handles exchanged once with ncclAllGather; signalling is one single-CTA
// kernel that writes a monotonically increasing epoch into each peer's flag slot and polls its own slots (bounded:
// a peer that never arrives raises an error flag instead of hanging the GPU).  Local back end: pointers are shared
// inside the process; signalling is an event + host barrier like the exchange below.
bool omg_comm_p2p(const omg_comm* c);
// collective: every rank contributes ONE device allocation (base pointer as returned by cudaMalloc); peer_ptrs[r]
// becomes a pointer to rank r's allocation valid on this rank's device (peer_ptrs[me] = ptr)
void omg_comm_share(omg_comm* c, void* ptr, void** peer_ptrs);
// collective: close every mapping made by omg_comm_share since the communicator was created (the flag arrays stay), then
// a barrier -- after it the owners may free the shared allocations (CUDA IPC: never free what a peer still has mapped)
void omg_comm_unshare(omg_comm* c);
// `epoch` (device, one counter per channel) is incremented when the wait completes: kernels launched after it see the
// new value, which selects the plane buffer the peers have just filled
// signal = 0: wait only (the puts of the current epoch were signalled from inside a kernel, see omg_fast.cu)
void omg_comm_signal_wait(omg_comm* c, int channel, const int* peers, int npeers, unsigned long long* epoch, int signal, cudaStream_t s);
// device pointers for in-kernel signalling: peer_flags[r] = rank r's flag array mapped here, flags = mine,
// both indexed [channel * nranks + writer rank]; err = bounded-wait error word
void omg_comm_p2p_ptrs(omg_comm* c, unsigned long long* const** peer_flags, unsigned long long** flags, int** err);
// non-zero once a bounded wait has expired: omg_p2p_err_code(peer, channel) of the first wait that gave up.  The word
// lives in mapped pinned host memory, so this is a plain host load (no device synchronisation).
int omg_comm_p2p_error(omg_comm* c);
// the bound of the in-kernel waits in ns (OMG_P2P_TIMEOUT_S, default 120 s; 0 = unbounded like MPI_Waitall)
unsigned long long omg_comm_p2p_timeout_ns(const omg_comm* c);
enum { OMG_P2P_CHANNELS = 64 };
__host__ __device__ inline int omg_p2p_err_code(int peer, int channel) { return 1 + peer + 64 * channel; }
inline int omg_p2p_err_peer(int code) { return (code - 1) % 64; }
inline int omg_p2p_err_channel(int code) { return (code - 1) / 64; }

// all ranks call these at the same program points with matching segments (throws std::runtime_error)
void omg_comm_exchange(omg_comm* c, const OmgXfer* segs, int nsegs, cudaStream_t s);
void omg_comm_allreduce(omg_comm* c, double* d_val, int op, cudaStream_t s);
// element-wise sum of n doubles over all ranks, in place (the mpi_allreduce of src/m_free_space.f90:152-153); the local
// back end adds the contributions in rank order
void omg_comm_allreduce_sum_vec(omg_comm* c, double* d_val, size_t n, cudaStream_t s);
