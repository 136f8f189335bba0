/* ffk_index.h — index-driven entry points: Embedding, Gather, TopK. Included by ffk.h.
 * Index tensors are INT32 or INT64 (ffk_dtype_t); rows/values selected by an index are
 * copied bit-exactly. */
#ifndef FFK_INDEX_H_
#define FFK_INDEX_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Aggregation codes: the reference's AggregateOp enum order (op-attrs/aggregate_op.enum.toml);
 * FFK_AGGR_NONE stands for std::nullopt. */
typedef enum { FFK_AGGR_NONE = -1, FFK_AGGR_SUM = 0, FFK_AGGR_AVG = 1 } ffk_aggregate_t;

/* Kernels::Embedding::forward_kernel — include/kernels/embedding_kernels.h:11,
 * src/cuda/embedding_kernels.cu:63-141,344-426.
 *   aggr NONE: output[b][:] = weight[input[b]][:]            (input: batch indices)
 *   aggr SUM : output[b][:] = sum_j weight[input[b][j]][:]    (input: batch x in_dim indices)
 *   aggr AVG : the mean over the bag (the reference rescales inside the loop,
 *              embedding_kernels.cu:106-114 — SURVEY App. A; the mean is implemented).
 * in_dim is the bag size (ignored for NONE), out_dim the row width. weight/output share
 * `dtype` (FLOAT, HALF, BFLOAT16, DOUBLE). Indices are not range-checked, as in the reference. */
int ffk_embedding_forward(void *stream, void const *input, void *output, void const *weight,
                          int index_dtype, int dtype, int aggr, int in_dim, int out_dim,
                          int batch_size);

/* Kernels::Embedding::backward_kernel — include/kernels/embedding_kernels.h:21,
 * src/cuda/embedding_kernels.cu:143-335,387-470. weight_grad[input[b][j]][:] += output_grad[b][:]
 * (x 1/in_dim for AVG) by floating-point atomics, as the reference; duplicate indices make
 * the summation order, and so the last bits, run-dependent. `output_grad` is the argument
 * the reference names `output`. */
int ffk_embedding_backward(void *stream, void const *input, void const *output_grad,
                           void *weight_grad, int index_dtype, int dtype, int aggr, int in_dim,
                           int out_dim, int batch_size);

/* Fused-operator form of Embedding backward + SGD for tables (additive): the
 * reference scatters dOut into a dense weight gradient (embedding_kernels.cu:144-335) and then
 * runs the dense optimizer over the whole table (optimizer_kernel.cu:20-41) — three passes over
 * a table of which a step touches a handful of rows. With plain SGD (no momentum, no weight decay,
 * which is how examples/cpp/DLRM trains its tables) W -= lr * scatter_add(dOut) is the same as
 * scattering -lr * dOut straight into W: only the looked-up rows move, no gradient table, no zero
 * fill, no dense pass. Same index / aggregation rules as ffk_embedding_backward. */
int ffk_embedding_sgd_update(void *stream, void const *input, void const *output_grad, void *weight,
                             float lr, int index_dtype, int dtype, int aggr, int in_dim, int out_dim,
                             int batch_size);

/* Kernels::Embedding::rand_generate_int{32,64}_wrapper — include/kernels/embedding_kernels.h:32-33,
 * src/cuda/embedding_kernels.cu:24-38,337-342: ptr[i] = i % p. */
int ffk_rand_generate_int(void *stream, void *ptr, size_t size, long long p, int index_dtype);

/* Kernels::Gather::forward_kernel — include/kernels/gather_kernels.h:19,
 * src/cuda/ops/gather_kernels.cu:26-50,121-147. torch.gather along one dimension with the
 * tensors viewed as input [outer][input_dim_size][stride], index/output
 * [outer][output_dim_size][stride]: output[o] = input[outer*stride*input_dim + index[o]*stride + rest]. */
int ffk_gather_forward(void *stream, void const *input, void const *index, void *output,
                       long long output_size, long long stride, long long input_dim_size,
                       long long output_dim_size, int index_dtype, int dtype);

/* Kernels::Gather::backward_kernel — include/kernels/gather_kernels.h:24,
 * src/cuda/ops/gather_kernels.cu:52-77,149-175. input_grad[...] += output_grad[o] (atomics). */
int ffk_gather_backward(void *stream, void const *output_grad, void const *index,
                        void *input_grad, long long output_size, long long stride,
                        long long input_dim_size, long long output_dim_size, int index_dtype,
                        int dtype);

/* Kernels::TopK::forward_kernel — include/kernels/topk_kernels.h:18,
 * src/cuda/ops/topk_kernels.cu:342-407. For each of batch_size rows of `length` floats: the k
 * largest values in descending order and their int32 positions; equal values keep the lower
 * index first (the TensorFlow rule the reference's heap implements). The output is sorted
 * whether or not `sorted` is set, which the reference's merge step also guarantees. */
int ffk_topk_forward(void *stream, float const *input, float *output, int *indices,
                     size_t batch_size, int length, int k, int sorted);

/* Kernels::TopK::backward_kernel — include/kernels/topk_kernels.h:27,
 * src/cuda/ops/topk_kernels.cu:410-437. in_grad[b][indices[b][i]] += out_grad[b][i]. */
int ffk_topk_backward(void *stream, float const *out_grad, int const *indices, float *in_grad,
                      size_t batch_size, int length, int k);

#ifdef __cplusplus
}
#endif
#endif
