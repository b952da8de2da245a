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
// all ranks call these at the same program points with matching segments (throws std::runtime_error)
void omg_comm_exchange(omg_comm* c, const OmgXfer* segs, int nsegs, cudaStream_t s);
void omg_comm_allreduce(omg_comm* c, double* d_val, int op, cudaStream_t s);
