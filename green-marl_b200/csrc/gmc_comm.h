// Internal: the communicator object behind gmc_comm (include/gm_cuda.h) and the collectives the sharded
// PageRank / BFS host code calls.  world == 1 (or a NULL communicator) makes every call a local no-op/copy.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>
#include <stdint.h>

struct gmc_comm {
    int world = 1, rank = 0;
    ncclComm_t nccl = nullptr;
    void* scratch = nullptr;
    bool local = false;     // every rank lives in THIS process (gmc_group): peers are reached with plain device pointers
    int device = -1;        // local ranks: the CUDA device of this rank
};

int gmc_comm_allgather(gmc_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st);
int gmc_comm_alltoall(gmc_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st);
int gmc_comm_allreduce_sum_f64(gmc_comm* c, const double* send, double* recv, size_t count, cudaStream_t st);
int gmc_comm_allreduce_sum_i64(gmc_comm* c, const long long* send, long long* recv, size_t count, cudaStream_t st);
int gmc_comm_allgather_segments(gmc_comm* c, void* buf, size_t elem_bytes, const int64_t* seg_off, const int64_t* seg_len,
                                int nseg, cudaStream_t st);
int gmc_comm_ipc_open_all(gmc_comm* c, void* local_base, void** peer_bases, cudaStream_t st);
// device-side barrier on `st` (a 1-element allreduce), no host synchronisation
int gmc_comm_stream_barrier(gmc_comm* c, cudaStream_t st);
void gmc_comm_ipc_close_all(int world, int rank, void** peer_bases, bool local);
