/* ffk_shape.h — layout and reduction entry points: Concat / Split (strided block copy),
 * Transpose, Reverse, Reduce, Pool2D, Conv2D. Included by ffk.h. */
#ifndef FFK_SHAPE_H_
#define FFK_SHAPE_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* copy_with_stride / add_with_stride — src/cuda/cuda_helper.cu:144-172, the kernels behind
 * Kernels::Concat (include/kernels/concat_kernels.h:11-19, src/cuda/ops/concat_kernels.cu:33-88)
 * and Kernels::Split (include/kernels/split_kernels.h:10-24, src/cuda/ops/split_kernels.cu:23-58):
 * for every block b < num_blocks and offset o < min(out_blk, in_blk):
 *   output[b*out_blk + o] (=|+=) input[b*in_blk + o].
 * Concat forward = one copy per input with output offset by the running block size; Split
 * forward = one copy per output with the input offset likewise; the backward forms use add. */
int ffk_copy_with_stride(void *stream, void *output, void const *input, long long num_blocks,
                         long long output_blk_size, long long input_blk_size, int dtype);
int ffk_add_with_stride(void *stream, void *output, void const *input, long long num_blocks,
                        long long output_blk_size, long long input_blk_size, int dtype);

/* Kernels::Transpose::forward_kernel / backward_kernel — include/kernels/transpose_kernels.h:23-31,
 * src/cuda/ops/transpose_kernels.cu:45-120. N-d permutation (ndims <= 5) in FF order with
 * torch.permute semantics: output.shape[i] = input.shape[perm[i]]. accumulate = 0 is the
 * forward (output = permuted input); the backward is the same call with the INVERSE
 * permutation and accumulate = 1 (in_grad += permuted out_grad). The reference's stride
 * computation adds 1 to every extent and its write is `out += out*beta + in`
 * (transpose_kernels.cu:58,75-78 — SURVEY App. A); the intended permutation is implemented. */
int ffk_transpose(void *stream, void const *input, void *output, int ndims,
                  long long const *input_dims_host, int const *perm_host, int accumulate,
                  int dtype);

/* Kernels::Reverse::forward_kernel / backward_kernel — include/kernels/reverse_kernels.h:10-24,
 * src/cuda/ops/reverse_kernels.cu:24-68. Tensor viewed as [num_out_blks][reverse_dim_size][in_blk_size];
 * out[b][r][i] = in[b][reverse_dim_size-1-r][i]. The backward is the same call on the gradients
 * (overwrite), as in the reference. (The reference kernel stores through its reduced index,
 * reverse_kernels.cu:29-36 — SURVEY App. A; the intended reversal is implemented.) */
int ffk_reverse(void *stream, void const *input, void *output, long long num_out_blks,
                long long reverse_dim_size, long long in_blk_size, int dtype);

/* Kernels::Reduce::forward_kernel — include/kernels/reduce_kernels.h:37,
 * src/cuda/ops/reduce_kernels.cu:47-65 (cudnnReduceTensor between the input and output
 * descriptors; the reference never configures the reduce descriptor — SURVEY App. A — so
 * op_type selects the intended REDUCE_SUM (48) or REDUCE_MEAN (45)). Shapes in FF order,
 * same rank; every dimension where out extent is 1 and in extent is not is reduced. */
int ffk_reduce_forward(void *stream, void const *input, void *output, int ndims,
                       long long const *input_dims_host, long long const *output_dims_host,
                       int op_type, int dtype);
/* Kernels::Reduce::backward_kernel — include/kernels/reduce_kernels.h:42,
 * src/cuda/ops/reduce_kernels.cu:67-92. input_grad += alpha * broadcast(output_grad),
 * alpha = 1 (SUM) or 1/reduction_size (MEAN). */
int ffk_reduce_backward(void *stream, void const *output_grad, void *input_grad, int ndims,
                        long long const *input_dims_host, long long const *output_dims_host,
                        int op_type, int dtype);

/* Pooling codes: the reference's PoolOp enum order (op-attrs/pool_op.enum.toml). */
typedef enum { FFK_POOL_MAX = 0, FFK_POOL_AVG = 1 } ffk_pool_t;

/* Kernels::Pool2D::forward_kernel — include/kernels/pool_2d_kernels.h:62,
 * src/cuda/ops/pool_2d_kernels.cu:94-110 (cudnnPoolingForward; MAX, or AVG with
 * COUNT_EXCLUDE_PADDING :61-66). NCHW fp32. */
int ffk_pool2d_forward(void *stream, float const *input, float *output, int n, int c, int in_h,
                       int in_w, int out_h, int out_w, int kernel_h, int kernel_w, int stride_h,
                       int stride_w, int pad_h, int pad_w, int pool_type);
/* Kernels::Pool2D::backward_kernel — include/kernels/pool_2d_kernels.h:67,
 * src/cuda/ops/pool_2d_kernels.cu:112-135 (cudnnPoolingBackward, beta = 1: input_grad
 * accumulates). MAX routes each output gradient to the first maximal element of its window. */
int ffk_pool2d_backward(void *stream, float const *input, float *input_grad, float const *output,
                        float const *output_grad, int n, int c, int in_h, int in_w, int out_h,
                        int out_w, int kernel_h, int kernel_w, int stride_h, int stride_w,
                        int pad_h, int pad_w, int pool_type);

/* Conv2D geometry: the integer arguments of Kernels::Conv2D::init_kernel plus the tensor
 * extents it reads from its accessors (include/kernels/conv_2d_kernels.h:47-59). NCHW fp32,
 * cross-correlation (cuDNN's CUDNN_CROSS_CORRELATION, the reference's mode), dilation 1. */
typedef struct {
  int n, in_c, in_h, in_w;
  int out_c, out_h, out_w;
  int kernel_h, kernel_w, pad_h, pad_w, stride_h, stride_w, groups;
} ffk_conv2d_config_t;

/* Scratch bytes for the im2col panel of ONE sample (the minimum); more lets several samples
 * share a GEMM launch. The reference passes handle.workSpace to cuDNN the same way
 * (conv_2d_kernels.cu:286-287). */
size_t ffk_conv2d_workspace_bytes(ffk_conv2d_config_t const *cfg_host);

/* Kernels::Conv2D::forward_kernel — include/kernels/conv_2d_kernels.h:61,
 * src/cuda/ops/conv_2d_kernels.cu:269-313: convolution, + bias (cudnnAddTensor) when bias is
 * not NULL, then ReLU when `relu` (cudnnActivationForward with the RELU descriptor). Filter
 * layout [out_c][in_c/groups][kernel_h][kernel_w]. */
int ffk_conv2d_forward(void *stream, ffk_conv2d_config_t const *cfg_host, float const *input,
                       float *output, float const *filter, float const *bias, int relu,
                       void *workspace, size_t workspace_bytes);

/* Kernels::Conv2D::backward_kernel — include/kernels/conv_2d_kernels.h:69,
 * src/cuda/ops/conv_2d_kernels.cu:315-387: optional reluBackward on output_grad in place,
 * then filter_grad, bias_grad and input_grad all ACCUMULATE (alpha = beta = 1). bias_grad and
 * input_grad may be NULL. */
int ffk_conv2d_backward(void *stream, ffk_conv2d_config_t const *cfg_host, float const *input,
                        float *input_grad, float const *output, float *output_grad,
                        float const *filter, float *filter_grad, float *bias_grad, int relu,
                        void *workspace, size_t workspace_bytes);

#ifdef __cplusplus
}
#endif
#endif
