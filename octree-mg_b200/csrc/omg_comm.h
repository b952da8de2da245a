// omg_comm.h -- inter-rank exchange (replaces sort_and_transfer_buffers, src/m_communication.f90:37-88)
#pragma once
#include <cuda_runtime.h>
struct omg_comm;
int omg_comm_unique_id(char out[128]);
void omg_comm_destroy(omg_comm* c);
bool omg_comm_graph_safe(omg_comm* c);
void omg_comm_halo(omg_comm* c, int lvl, cudaStream_t s);
void omg_comm_restrict(omg_comm* c, int lvl, cudaStream_t s);
void omg_comm_prolong(omg_comm* c, int lvl, cudaStream_t s);
void omg_comm_allreduce_max(omg_comm* c, double* d_val, cudaStream_t s);
void omg_comm_allreduce_sum(omg_comm* c, double* d_val, cudaStream_t s);
