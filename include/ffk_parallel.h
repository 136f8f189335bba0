/* ffk_parallel.h — NCCL entry points of the kernel API: communicator bootstrap and the
 * gradient-synchronising optimizer updates. Included by ffk.h.
 * NCCL is resolved at run time (dlopen of libnccl.so.2): libffk.so itself does not link it, and
 * a process that already carries NCCL (e.g. through torch) shares that copy. The single-device
 * forms of Replicate / Reduction / Repartition / Combine are in ffk_elementwise.h; their
 * multi-GPU forms are composed from these collectives by the host layer
 * (flexflow-train_b200/parallel_ops.py). */
#ifndef FFK_PARALLEL_H_
#define FFK_PARALLEL_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

#define FFK_NCCL_UNIQUE_ID_BYTES 128

/* Kernels::NCCL::generate_unique_id — include/kernels/nccl.h:30, src/cuda/nccl.cc:8-15
 * (ncclGetUniqueId). Writes FFK_NCCL_UNIQUE_ID_BYTES to host memory; rank 0 calls it and ships
 * the bytes to the other ranks by any means. */
int ffk_nccl_get_unique_id(void *id_out_host);

/* Kernels::NCCL::create_comm — include/kernels/nccl.h:31, src/cuda/nccl.cc:17-29
 * (ncclCommInitRank on the current device). The result is what PerDeviceFFHandle::ncclComm holds. */
int ffk_nccl_comm_create(void **comm_out, void const *id_host, int num_ranks, int my_rank);
int ffk_nccl_comm_destroy(void *comm);

/* In-place sum over ranks (ncclAllReduce, ncclSum — the call at src/cuda/optimizer_kernel.cu:83-89) of
 * `count` elements: the exchange step of Replicate-backward / Reduction-forward
 * (replicate_kernels.cu:66-78, reduction_kernels.cu:67-79 sum stacked replicas on one device) and of
 * the *_nccl_update tasks. FLOAT, HALF, BFLOAT16,
 * DOUBLE, INT32, INT64. */
int ffk_all_reduce_sum(void *stream, void *comm, void *buffer, size_t count, int dtype);

/* sgd_nccl_update_task_gpu — include/kernels/optimizer_kernels.h:19 (its missing comma fixed),
 * src/cuda/optimizer_kernel.cu:73-104: ncclAllReduce(weight_grad, SUM — no division by the
 * replica count) on `comm`, then the SGD update of ffk_sgd_update. */
int ffk_sgd_nccl_update(void *stream, void *comm, float lr, float momentum, int nesterov,
                        float weight_decay, void *weight_grad, size_t size, void *weight,
                        float *sgd_v, int dtype, float *master_weight);

/* adam_nccl_update_task_gpu — include/kernels/optimizer_kernels.h:41,
 * src/cuda/optimizer_kernel.cu:181-214. */
int ffk_adam_nccl_update(void *stream, void *comm, float alpha_t, float beta1, float beta2,
                         float weight_decay, float epsilon, void *weight_grad, size_t size,
                         float *adam_m, float *adam_v, void *weight, int dtype,
                         float *master_weight);

/* ---- Sum over ranks through NVLink peer memory (no counterpart in the reference, which only
 * has the NCCL call above, src/cuda/optimizer_kernel.cu:83-89; same SUM-without-division contract). One process per GPU on one node. Each rank creates a region of
 * `data_bytes` (device memory, cudaMalloc) and gets a 64-byte handle; the handles of all ranks,
 * concatenated in rank order and shipped by any means, connect the group (cudaIpcOpenMemHandle).
 * Tensors that take part in the collective live inside the region (ffk_p2p_local_buffer). */
#define FFK_P2P_HANDLE_BYTES 64
int ffk_p2p_create(void **group_out, int rank, int world, size_t data_bytes,
                   unsigned char *handle_out /* FFK_P2P_HANDLE_BYTES */);
int ffk_p2p_connect(void *group, unsigned char const *handles_all /* world * FFK_P2P_HANDLE_BYTES */);
int ffk_p2p_local_buffer(void *group, void **data_out, size_t *bytes_out);

/* out[i] = sum over ranks of in[i] for `count` elements — what ncclAllReduce(ncclSum) gives the
 * *_nccl_update tasks (src/cuda/optimizer_kernel.cu:83-89, 191-197) — (FLOAT, HALF, BFLOAT16; fp32
 * accumulation in rank order, so every rank ends with bit-identical results). `in_offset` and
 * `out_offset` are byte offsets into the region, the same on every rank; the two ranges must not
 * overlap; offsets and size are multiples of 16 bytes. Up to 64 KB every rank pushes its input
 * into receive slots of all peers and sums what arrived (one switch traversal); up to 512 KB
 * every rank reads all peers (one shot); above, rank r reduces chunk r and the chunks are
 * gathered (two shots). Every rank
 * must issue the group's collectives in the same order. The input may be overwritten and the
 * output read as soon as the kernel has finished on the caller's stream. Waits for peers are
 * bounded (30 s; FFK_P2P_TIMEOUT_MS, read by ffk_p2p_create, overrides). A wait that expires
 * never publishes a sum: the failure flag is raised on EVERY rank of the group, each rank's
 * output range of the failed launch and of every later launch is filled with NaN bit patterns,
 * and ffk_p2p_status returns FFK_ERR_CUDA on all ranks until the group is destroyed. */
int ffk_p2p_all_reduce_sum(void *stream, void *group, size_t in_offset, size_t out_offset,
                           size_t count, int dtype);
/* The same sum (same rank-order fp32 arithmetic, bit-identical results) by a kernel shaped to
 * run UNDERNEATH other work: `blocks` CTAs of 128 threads, at most 80 registers per thread and
 * no shared memory, so that a CTA fits on an SM beside a resident tcgen05 GEMM CTA. Meant for
 * the data-parallel gradient sum of optimizer_kernel.cu:83-89 launched on a side stream while
 * the backward pass is still producing the next bucket; the all-reduce of a layer's gradients
 * then costs no step time as long as NVLink keeps up (32 blocks move ~150 GB/s). Every rank
 * must pass the same `blocks`. */
int ffk_p2p_all_reduce_sum_background(void *stream, void *group, size_t in_offset,
                                      size_t out_offset, size_t count, int dtype, int blocks);
int ffk_p2p_status(void *group);
int ffk_p2p_destroy(void *group);

#ifdef __cplusplus
}
#endif
#endif
