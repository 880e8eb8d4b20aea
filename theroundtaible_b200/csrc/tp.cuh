// Tensor parallelism plumbing: NCCL loaded lazily (dlopen) so that single-GPU users need no NCCL at all.
// One engine per GPU/process; the communicator spans the tp_size engines.  No reference counterpart
// (/root/reference has no collective call site); SURVEY.md 8(e) defines the partitioning:
// QKV / gate / up column-parallel, O / down row-parallel + all-reduce, KV heads / TP, lm_head vocab-parallel.
#pragma once
#include "common.cuh"

struct TpComm;                                     // opaque
void tp_unique_id(void* out128);
TpComm* tp_init(int rank, int size, const void* id128);
void tp_destroy(TpComm* c);
void tp_allreduce_sum_f32(TpComm* c, float* buf, size_t count, cudaStream_t st);
void tp_allreduce_sum_f16(TpComm* c, void* buf, size_t count, cudaStream_t st);
void tp_allgather_bytes(TpComm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st);
// `rows` all-gathers as ONE grouped NCCL launch: row r gathers bytes_per_rank from send + r*send_stride into recv + r*recv_stride
void tp_allgather_rows(TpComm* c, const void* send, size_t send_stride, void* recv, size_t recv_stride, size_t bytes_per_rank, int rows,
                       cudaStream_t st);

// Exchange a cudaMalloc'ed buffer with every rank through CUDA IPC: out_ptrs[p] is rank p's buffer mapped into this
// process (out_ptrs[rank] = local).  Collective: every rank must call it in the same order.
void tp_ipc_exchange(TpComm* c, void* local, void** out_ptrs, cudaStream_t st);
void tp_ipc_close(TpComm* c, void** ptrs);
int tp_rank(const TpComm* c);
int tp_size(const TpComm* c);
