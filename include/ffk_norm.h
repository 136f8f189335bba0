/* ffk_norm.h — HBM-bound row/channel normalisations: Softmax, LayerNorm, BatchNorm.
 * Included by ffk.h. */
#ifndef FFK_NORM_H_
#define FFK_NORM_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Kernels::Softmax::forward_kernel — include/kernels/softmax_kernels.h:27,
 * src/cuda/ops/softmax_kernels.cu:45-61 (cudnnSoftmaxForward, ACCURATE, MODE_CHANNEL over the
 * (n,c,h,w) given to init_kernel). Here the tensor is viewed as [outer][dim][inner] and the
 * softmax runs over `dim`; the reference's case is outer = n, dim = c, inner = h*w. */
int ffk_softmax_forward(void *stream, void const *input, void *output, long long outer,
                        long long dim, long long inner, int dtype);

/* Kernels::Softmax::backward_kernel — include/kernels/softmax_kernels.h:32,
 * src/cuda/ops/softmax_kernels.cu:63-73. mode 0 is the reference: input_grad <- output_grad
 * (a device copy; only right when the op is fused with a cross-entropy loss).
 * mode 1 is the full Jacobian product dx = y (.) (dy - sum(dy (.) y)) and needs `output`. */
int ffk_softmax_backward(void *stream, void *input_grad, void const *output_grad,
                         void const *output, long long outer, long long dim, long long inner,
                         int dtype, int mode);

/* Kernels::LayerNorm::forward_kernel — include/kernels/layer_norm_kernels.h:44,
 * src/cuda/ops/layer_norm_kernels.cu:68-112,325-350.
 *   mean[i], rstd[i] over the N = effective_num_elements of row i (var = E[x^2]-mean^2
 *   clamped at 0, as the reference), y = (x - mean) * rstd * gamma + beta.
 * gamma/beta may be NULL (no affine). mean/rstd are fp32 [M] and are kept for backward:
 * they are LayerNormPerDeviceState::mean / ::rstd. */
int ffk_layer_norm_forward(void *stream, void const *input, void *output, void const *gamma,
                           void const *beta, float *mean, float *rstd, long long M, long long N,
                           float eps, int dtype);

/* Kernels::LayerNorm::backward_kernel — include/kernels/layer_norm_kernels.h:51,
 * src/cuda/ops/layer_norm_kernels.cu:114-289,352-408.
 *   input_grad += rstd*dY*gamma + c1*X + c2  (the dX formula of layer_norm_kernels.cu:163-181,
 *     which the reference defines but never launches — SURVEY App. A; NULL skips it);
 *   gamma_grad[j] = sum_i dY[i][j] (X[i][j]-mean[i]) rstd[i];  beta_grad[j] = sum_i dY[i][j]
 *     (overwritten like the reference when accumulate_param_grads == 0, += otherwise).
 * `workspace` (device, optional) lets one pass over dY and X produce all three results:
 * it holds per-CTA partial column sums, ffk_layer_norm_backward_workspace_bytes(N) bytes for
 * full speed, any smaller multiple of 8*N bytes for fewer CTAs; with workspace == NULL the
 * parameter gradients are computed by a second pass over dY and X. */
size_t ffk_layer_norm_backward_workspace_bytes(long long N);
int ffk_layer_norm_backward(void *stream, void const *output_grad, void const *input,
                            void *input_grad, void const *gamma, void *gamma_grad,
                            void *beta_grad, float const *mean, float const *rstd, long long M,
                            long long N, int dtype, int accumulate_param_grads, void *workspace,
                            size_t workspace_bytes);

/* ffk_layer_norm_backward (layer_norm_kernels.h:47, layer_norm_kernels.cu:410-437) with the choice the
 * reference's convention does not offer:
 * accumulate_input_grad == 0 WRITES input_grad instead of adding to it (one HBM pass less, and no
 * zero fill before it) for callers that know this op is the first writer of that gradient —
 * the step driver's fused-operator execution. 1 = ffk_layer_norm_backward. */
int ffk_layer_norm_backward_ex(void *stream, void const *output_grad, void const *input,
                               void *input_grad, void const *gamma, void *gamma_grad,
                               void *beta_grad, float const *mean, float const *rstd, long long M,
                               long long N, int dtype, int accumulate_param_grads,
                               int accumulate_input_grad, void *workspace, size_t workspace_bytes);

/* No counterpart in the reference. LayerNorm backward for callers that cannot allocate device
 * memory and have no workspace to pass: tasks running under a runtime that has already claimed
 * the GPU's memory for its own pool (the reference's legacy Legion/Realm runtime does so at
 * start-up, so a later cudaMalloc may fail, and in any case synchronises the device), or code
 * that reaches the backward call without an Allocator / PerDeviceFFHandle workspace. Same
 * arguments and results as ffk_layer_norm_backward without the workspace pair: the partial
 * dgamma / dbeta sums are kept in a 4 MiB array built into libffk.so (held by every GPU that
 * loads the library, until exit). Limit: N <= 16384, else FFK_ERR_LIMIT. It allocates nothing.
 * The first call on a device looks up the array's device address — make that call before any
 * stream capture. Calls on one GPU share the array: run them one after another (same stream
 * or event-ordered); calls that overlap on the GPU race. */
int ffk_layer_norm_backward_static(void *stream, void const *output_grad, void const *input,
                                   void *input_grad, void const *gamma, void *gamma_grad,
                                   void *beta_grad, float const *mean, float const *rstd,
                                   long long M, long long N, int dtype,
                                   int accumulate_param_grads);

/* Kernels::BatchNorm::forward_kernel — include/kernels/batch_norm_kernels.h:49,
 * src/cuda/ops/batch_norm_kernels.cu:26-53 (cudnnBatchNormalizationForwardTraining, SPATIAL,
 * exponentialAverageFactor 1.0). NCHW fp32. Per channel: batch mean and biased variance,
 * y = scale*(x-mean)/sqrt(var+eps)+bias; save_mean/save_invstd keep mean and 1/sqrt(var+eps)
 * for backward; running_mean/var = factor*batch + (1-factor)*running with the UNBIASED
 * variance, cuDNN's rule. eps: the reference passes CUDNN_BN_MIN_EPSILON (0.0 in cuDNN 9). */
int ffk_batch_norm_forward(void *stream, float const *input, float *output, float const *scale,
                           float const *bias, float *running_mean, float *running_var,
                           float *save_mean, float *save_invstd, int n, int c, int h, int w,
                           float eps, float factor);

/* Kernels::BatchNorm::backward_kernel — include/kernels/batch_norm_kernels.h:58,
 * src/cuda/ops/batch_norm_kernels.cu:55-90. If relu != 0 first dY <- (Y > 0) ? dY : 0 in place
 * (reluBackward), then cudnnBatchNormalizationBackward with every alpha/beta = 1:
 * input_grad, scale_grad and bias_grad all accumulate. */
int ffk_batch_norm_backward(void *stream, float const *input, float *output_grad,
                            float const *output, float *input_grad, float const *scale,
                            float *scale_grad, float *bias_grad, float const *save_mean,
                            float const *save_invstd, int n, int c, int h, int w, int relu);

#ifdef __cplusplus
}
#endif
#endif
