// Compile-time check (built by __graft_entry__.build(), never run): the veneer's metrics wrappers,
// PerfMetrics helpers and initialiser functions exist with the reference's names and signatures
// (include/kernels/metrics_kernels.h:8-21, perf_metrics.h:36-40, initializer_kernels.h:21-28, cpu.h:6).
#include "kernels/cpu.h"
#include "kernels/initializer_kernels.h"
#include "kernels/metrics_kernels.h"
#include "kernels/perf_metrics.h"

using namespace FlexFlow;

void (*const p_sparse)(ffStream_t, MetricsAttrs const &, float const *, int const *, int, int, PerfMetrics &) =
    &update_metrics_sparse_label_kernel;
void (*const p_label)(ffStream_t, MetricsAttrs const &, float const *, float const *, int, int, PerfMetrics &) =
    &update_metrics_label_kernel;
PerfMetrics (*const p_update)(PerfMetrics const &, PerfMetrics const &) = &update;
PerfMetrics (*const p_scale)(PerfMetrics const &, float) = &apply_scale;
void (*const p_zero)(TaskLocation const &, GenericTensorAccessorW const &) = &zero_init_kernel;
void (*const p_zero_gpu)(GenericTensorAccessorW const &) = &zero_init_kernel_gpu;
void (*const p_zero_cpu)(GenericTensorAccessorW const &) = &zero_init_kernel_cpu;
void (*const p_const)(TaskLocation const &, GenericTensorAccessorW const &, DataTypeValue) = &constant_init_kernel;
void (*const p_const_gpu)(GenericTensorAccessorW const &, DataTypeValue) = &constant_init_kernel_gpu;
void (*const p_const_cpu)(GenericTensorAccessorW const &, DataTypeValue) = &constant_init_kernel_cpu;

int host_side_semantics() {
  // the host-only pieces can be exercised without a device
  MetricsAttrs m(LossFunction::SPARSE_CATEGORICAL_CROSSENTROPY, {Metric::ACCURACY, Metric::SPARSE_CATEGORICAL_CROSSENTROPY});
  if (ffk_metric_flags(m) != (FFK_METRIC_ACCURACY | FFK_METRIC_SPARSE_CCE)) return 1;
  PerfMetrics a(0.0), b(0.0);
  a.train_all = 3; a.train_correct = 2; a.sparse_cce_loss = 1.5f;
  b.train_all = 5; b.train_correct = 1; b.sparse_cce_loss = 0.5f; b.current_time = 7.0;
  PerfMetrics c = update(a, b);
  if (c.train_all != 8 || c.train_correct.value() != 3 || c.sparse_cce_loss.value() != 2.0f || c.current_time != 7.0) return 2;
  if (c.cce_loss.has_value()) return 3;                       // not measured stays empty
  if (apply_scale(c, 0.5f).sparse_cce_loss.value() != 1.0f) return 4;
  float host[6] = {1, 2, 3, 4, 5, 6};
  GenericTensorAccessorW w{{DataType::FLOAT, ArrayShape(std::vector<size_t>{2, 3}), host}};
  constant_init_kernel(TaskLocation::CPU, w, DataTypeValue(2.5f));
  for (float v : host) if (v != 2.5f) return 5;
  zero_init_kernel(TaskLocation::CPU, w);
  for (float v : host) if (v != 0.f) return 6;
  return 0;
}

#ifdef VENEER_EXTRAS_MAIN
int main() { return host_side_semantics(); }
#endif
