/* ffk_train.h — loss gradients, optimizer updates and metric accumulation. Included by ffk.h.
 * The reference's bodies for these live in files its own build does not compile
 * (src/cuda/loss_function_kernels.cu, optimizer_kernel.cu, metrics_functions.cu — SURVEY F5);
 * their headers are the API and their kernels the statement of the math. */
#ifndef FFK_TRAIN_H_
#define FFK_TRAIN_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* sparse_categorical_crossentropy_loss_backward_kernel — include/kernels/loss_function_kernels.h:8,
 * src/cuda/loss_function_kernels.cu:21-31,59-84.
 *   logit_grad = logit; logit_grad[i][label[i / k]] -= 1 for i < num_samples; logit_grad *= scale_factor * k
 * in ONE pass on the caller's stream (the reference uses a blocking cudaMemcpy and two more
 * kernels on a stream of its own). logit is [num_samples][num_classes]; label holds
 * num_samples / k int32 class ids. */
int ffk_sparse_categorical_crossentropy_loss_backward(void *stream, void *logit_grad,
                                                      void const *logit, int const *label,
                                                      int num_samples, int num_classes, int k,
                                                      float scale_factor, int dtype);

/* categorical_crossentropy_loss_backward_kernel — loss_function_kernels.h:18,
 * loss_function_kernels.cu:33-40,86-103: logit_grad = (logit - label) * scale_factor.
 * mean_squared_error_avg_loss_backward_kernel — loss_function_kernels.h:25,
 * loss_function_kernels.cu:42-49,105-122: the same expression. */
int ffk_categorical_crossentropy_loss_backward(void *stream, void *logit_grad, void const *logit,
                                               void const *label, size_t volume,
                                               float scale_factor, int dtype);
int ffk_mean_squared_error_avg_loss_backward(void *stream, void *logit_grad, void const *logit,
                                             void const *label, size_t volume,
                                             float scale_factor, int dtype);

/* identity_loss_backward_kernel — loss_function_kernels.h:32,
 * loss_function_kernels.cu:51-57,124-139: loss_grad = 1 * scale_factor. */
int ffk_identity_loss_backward(void *stream, void *loss_grad, void const *loss, size_t volume,
                               float scale_factor, int dtype);

/* sgd_ps_update_task_gpu — include/kernels/optimizer_kernels.h:8,
 * src/cuda/optimizer_kernel.cu:20-71. If num_replicas > 1 the stacked replica gradients
 * grad[j*size + i] are first summed in order j = 0.. (the reference folds them into replica 0
 * in place; here the sum stays in registers and `weight_grad` is not modified). Then
 *   g = dW + weight_decay*W; if momentum > 0 { V = momentum*V + g; g = nesterov ? g + momentum*V : V }
 *   W -= lr * g.
 * `dtype` is the dtype of weight and weight_grad; V is always fp32. `master_weight` (fp32,
 * optional) is the extension for 16-bit training: when given, the update is applied to it
 * and `weight` receives the rounded copy. */
int ffk_sgd_update(void *stream, float lr, float momentum, int nesterov, float weight_decay,
                   void const *weight_grad, size_t size, int num_replicas, void *weight,
                   float *sgd_v, int dtype, float *master_weight);

/* adam_ps_update_task_gpu — include/kernels/optimizer_kernels.h:30,
 * src/cuda/optimizer_kernel.cu:123-179.
 *   g = dW + weight_decay*W; M = b1*M + (1-b1)*g; V = b2*V + (1-b2)*g*g; W -= alpha_t*M/(sqrt(V)+eps)
 * with alpha_t computed by the caller (lib/runtime/src/optimizer.cc:140-146). Replica
 * summation and master weights as for SGD. */
int ffk_adam_update(void *stream, float alpha_t, float beta1, float beta2, float weight_decay,
                    float epsilon, void const *weight_grad, size_t size, int num_replicas,
                    float *adam_m, float *adam_v, void *weight, int dtype, float *master_weight);

/* Device-side PerfMetrics accumulator: the numeric fields of the reference's PerfMetrics
 * (include/kernels/perf_metrics.h:24-33) in a plain struct the kernels add into. */
typedef struct {
  int train_all;
  int train_correct;
  float cce_loss;
  float sparse_cce_loss;
  float mse_loss;
  float rmse_loss;
  float mae_loss;
} ffk_perf_metrics_t;

/* Bit flags for the `metrics` argument: the booleans of the reference's Metrics object
 * (src/cuda/metrics_functions.cu:23-71 reads metrics.measure_*). */
typedef enum {
  FFK_METRIC_ACCURACY = 1,
  FFK_METRIC_CCE = 2,
  FFK_METRIC_SPARSE_CCE = 4,
  FFK_METRIC_MSE = 8,
  FFK_METRIC_RMSE = 16,
  FFK_METRIC_MAE = 32
} ffk_metric_flag_t;

/* update_metrics_sparse_label_kernel — include/kernels/metrics_kernels.h:8,
 * src/cuda/metrics_functions.cu:23-71. Adds this batch's counts / losses into *perf (device
 * memory). Per-sample rules as the reference (arg-max starts from -1.0, log clamps at 1e-8). */
int ffk_update_metrics_sparse_label(void *stream, int metrics, float const *logit,
                                    int const *label, int num_samples, int num_classes,
                                    ffk_perf_metrics_t *perf);

/* update_metrics_label_kernel — include/kernels/metrics_kernels.h:15,
 * src/cuda/metrics_functions.cu:73-137 (dense one-hot / regression labels). */
int ffk_update_metrics_label(void *stream, int metrics, float const *logit, float const *label,
                             int num_samples, int num_classes, ffk_perf_metrics_t *perf);

#ifdef __cplusplus
}
#endif
#endif
