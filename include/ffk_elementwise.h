/* ffk_elementwise.h — HBM-bound pointwise entry points: ElementUnary, ElementBinary, Cast,
 * Dropout and the copy/accumulate helpers. Included by ffk.h.
 *
 * `op_type` arguments carry values of the reference's OperatorType enum
 * (lib/op-attrs/include/op-attrs/operator_type.enum.toml, 0-based in file order); the
 * ones these kernels accept are listed here. */
#ifndef FFK_ELEMENTWISE_H_
#define FFK_ELEMENTWISE_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  FFK_OP_SCALAR_MULTIPLY = 8,
  FFK_OP_SCALAR_ADD = 9,
  FFK_OP_SCALAR_TRUE_DIV = 11,
  FFK_OP_SCALAR_SUB = 12,
  FFK_OP_RELU = 13,
  FFK_OP_IDENTITY = 14,
  FFK_OP_SIGMOID = 15,
  FFK_OP_TANH = 16,
  FFK_OP_ELU = 17,
  FFK_OP_EW_ADD = 28,
  FFK_OP_EW_MUL = 29,
  FFK_OP_EW_SUB = 35,
  FFK_OP_EW_DIV = 36,
  FFK_OP_EW_MAX = 40,
  FFK_OP_EW_MIN = 41,
  FFK_OP_REDUCE_MEAN = 45,
  FFK_OP_REDUCE_SUM = 48,
  FFK_OP_EXP = 56,
  FFK_OP_SIN = 61,
  FFK_OP_COS = 62,
  FFK_OP_GELU = 67,
  FFK_OP_RSQRT = 70,
  FFK_OP_POW = 71
} ffk_op_type_t;

/* Kernels::ElementUnary::forward_kernel — include/kernels/element_unary_kernels.h:29,
 * src/cuda/ops/element_unary_kernels.cu:98-165,339-355. y = f(x) over n elements.
 * `scalar` is attrs.scalar for SCALAR_* / POW and the ELU coefficient (the reference's
 * cuDNN descriptor uses 0.0, element_unary_kernels.cu:81). GELU is the erf form. */
int ffk_element_unary_forward(void *stream, int op_type, float scalar, void const *input,
                              void *output, size_t n, int dtype);

/* Kernels::ElementUnary::backward_kernel — include/kernels/element_unary_kernels.h:36,
 * src/cuda/ops/element_unary_kernels.cu:167-245,357-373. input_grad += dy * f'(.) for every
 * op (the reference overwrites for POW/GELU/RSQRT — SURVEY App. A; accumulate chosen), and
 * GELU uses the true derivative Phi(x) + x phi(x) (the reference's expression has a sign and
 * constant slip, element_unary_kernels.cu:222-228). */
int ffk_element_unary_backward(void *stream, int op_type, float scalar, void const *input,
                               void *input_grad, void const *output, void const *output_grad,
                               size_t n, int dtype);

/* Kernels::ElementBinary::forward_kernel — include/kernels/element_binary_kernels.h:40,
 * src/cuda/ops/element_binary_kernels.cu:141-243. out = lhs (op) rhs with op in
 * {EW_ADD, EW_SUB, EW_MUL, EW_DIV, EW_MAX, EW_MIN}. Shapes are given in FF order
 * (outermost first) as host arrays of `ndims` extents; a side broadcasts along every
 * dimension where its extent is 1 and the output's is not. ndims == 0 means "all three
 * tensors have n = out_dims_host[0] elements of identical shape". */
int ffk_element_binary_forward(void *stream, int op_type, void const *lhs, void const *rhs,
                               void *out, int ndims, long long const *lhs_dims_host,
                               long long const *rhs_dims_host, long long const *out_dims_host,
                               int dtype);

/* Kernels::ElementBinary::backward_kernel — include/kernels/element_binary_kernels.h:49,
 * src/cuda/ops/element_binary_kernels.cu:26-80,245-422. lhs_grad / rhs_grad accumulate;
 * either may be NULL; a broadcast side receives the sum over its broadcast dimensions. */
int ffk_element_binary_backward(void *stream, int op_type, void const *out_grad, void const *lhs,
                                void const *rhs, void *lhs_grad, void *rhs_grad, int ndims,
                                long long const *lhs_dims_host, long long const *rhs_dims_host,
                                long long const *out_dims_host, int dtype);

/* Kernels::Cast::forward_kernel — include/kernels/cast_kernels.h:14,
 * src/cuda/ops/cast_kernels.cu:25-31,40-60. out[i] = (out_dtype) in[i]. */
int ffk_cast_forward(void *stream, void const *input, void *output, size_t n, int in_dtype,
                     int out_dtype);
/* Kernels::Cast::backward_kernel — include/kernels/cast_kernels.h:20,
 * src/cuda/ops/cast_kernels.cu:33-38,62-77. dst[i] = (dst_dtype) src[i] + dst[i]. */
int ffk_cast_backward(void *stream, void const *src, void *dst, size_t n, int src_dtype,
                      int dst_dtype);

/* Device-to-device copy and accumulate of n elements: the whole body of Flat / Reshape /
 * Combine / Repartition (single-device form) / Replicate-forward / Reduction-backward /
 * Softmax-backward. copy is bit-exact for every dtype.
 * References: src/cuda/ops/flat_kernels.cu:28,41; reshape_kernels.cu:34,48;
 * combine_kernels.cu:30,44; partition_kernels.cu:30,44; replicate_kernels.cu:42;
 * reduction_kernels.cu:59; softmax_kernels.cu:68; cuda_helper.cu:52,138. */
int ffk_copy(void *stream, void *dst, void const *src, size_t n, int dtype);
int ffk_accumulate(void *stream, void *dst, void const *src, size_t n, int dtype);

/* Sum of `num_replicas` stacked slabs: dst[i] (+)= sum_j src[i + j*n].
 * Replicate::backward_kernel (accumulate = 1; include/kernels/replicate_kernels.h:15,
 * src/cuda/ops/replicate_kernels.cu:25-34,50-62) and Reduction::forward_kernel
 * (accumulate = 0; include/kernels/reduction_kernels.h:11, src/cuda/ops/reduction_kernels.cu:25-50).
 * Slabs are added in index order j = 0,1,... in fp32, so results are reproducible. */
int ffk_sum_replicas(void *stream, void *dst, void const *src, size_t n, int num_replicas,
                     int accumulate, int dtype);

/* Kernels::Dropout — include/kernels/dropout_kernels.h:37-57, src/cuda/ops/dropout_kernels.cu:24-105.
 * y = x * keep / (1 - rate), keep ~ Bernoulli(1 - rate) from a Philox4x32-10 stream (16 bits per element) keyed by
 * (seed, offset); the keep decisions are written as one bit per element to `mask`
 * (ceil(n/32) uint32 words — the analogue of cuDNN's reserve space) and reused by backward,
 * which OVERWRITES input_grad like the reference does. */
size_t ffk_dropout_mask_bytes(size_t n);
int ffk_dropout_forward(void *stream, void const *input, void *output, void *mask, size_t n,
                        float rate, unsigned long long seed, unsigned long long offset,
                        int dtype);
int ffk_dropout_backward(void *stream, void const *output_grad, void *input_grad,
                         void const *mask, size_t n, float rate, int dtype);

#ifdef __cplusplus
}
#endif
#endif
