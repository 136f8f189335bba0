// Compile-time check (built by __graft_entry__.build(), never run): the veneer's metrics wrappers,
// PerfMetrics helpers and initialiser functions exist with the reference's names and signatures
// (include/kernels/metrics_kernels.h:8-21, perf_metrics.h:36-40, initializer_kernels.h:21-28, cpu.h:6).
#include "kernels/cpu.h"
#include "kernels/initializer_kernels.h"
#include "kernels/metrics_kernels.h"
#include "kernels/legion_dim_t.dtg.h"
#include "kernels/perf_metrics.h"
#include "kernels/profiling_settings.dtg.h"

#include <unordered_set>

#include <cmath>

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

// accessor.h:68-167, datatype_dispatch.h, device.h:110-114
int32_t *(*const p_i32)(GenericTensorAccessorW const &) = &get_int32_ptr;
double const *(*const p_f64)(GenericTensorAccessorR const &) = &get_double_ptr;
std::vector<float *> (*const p_fptrs)(std::vector<GenericTensorAccessorW> const &) = &get_float_ptrs;
std::vector<int64_t const *> (*const p_i64ptrs)(std::vector<GenericTensorAccessorR> const &) = &get_int64_ptrs;
std::pair<ArrayShape, DataType> (*const p_sd)(GenericTensorAccessorR const &) = &get_shape_and_datatype;
ffError_t (*const p_ev)(float *, ffEvent_t &, ffEvent_t &) = &ffEventElapsedTime;

template <DataType DT> struct SizeOf {
  size_t operator()(size_t n) const { return n * sizeof(real_type_t<DT>); }
};
template <DataType IT, DataType OT> struct Widen {
  void operator()(int *out) const { *out = int(sizeof(real_type_t<OT>)) - int(sizeof(real_type_t<IT>)); }
};

void (*const p_uni)(GenericTensorAccessorW const &, size_t, float, float) = &uniform_init_kernel;
void (*const p_glo)(GenericTensorAccessorW const &, size_t, float) = &glorot_uniform_init_kernel;
void (*const p_nrm)(GenericTensorAccessorW const &, size_t, float, float) = &norm_init_kernel;

int random_initialisers() {
  size_t const n = 20001;   // odd: the Box-Muller tail
  std::vector<float> a(n), b(n), c(n);
  auto acc = [&](std::vector<float> &v) {
    return GenericTensorAccessorW{{DataType::FLOAT, ArrayShape(std::vector<size_t>{v.size()}), v.data()}};
  };
  uniform_init_kernel(acc(a), 7, -0.25f, 0.75f);
  uniform_init_kernel(acc(b), 7, -0.25f, 0.75f);
  uniform_init_kernel(acc(c), 8, -0.25f, 0.75f);
  if (a != b || a == c) return 20;                       // a function of the seed
  double sum = 0;
  for (float v : a) {
    if (!(v >= -0.25f && v < 0.75f)) return 21;
    sum += v;
  }
  if (std::fabs(sum / n - 0.25) > 0.01) return 22;
  glorot_uniform_init_kernel(acc(b), 7, 0.5f);
  for (size_t i = 0; i < n; ++i)
    if (std::fabs(b[i] - (a[i] + 0.25f - 0.5f)) > 1e-6f) return 23;   // same stream, range [-0.5, 0.5)
  norm_init_kernel(acc(c), 11, 2.0f, 0.5f);
  double m = 0, v2 = 0;
  for (float v : c) m += v;
  m /= n;
  for (float v : c) v2 += (v - m) * (v - m);
  if (std::fabs(m - 2.0) > 0.02 || std::fabs(std::sqrt(v2 / n) - 0.5) > 0.02) return 24;
  for (float v : c) if (!std::isfinite(v)) return 25;
  int32_t ints[4] = {0, 0, 0, 0};
  GenericTensorAccessorW wi{{DataType::INT32, ArrayShape(std::vector<size_t>{4}), ints}};
  try {
    uniform_init_kernel(wi, 1, 0.f, 1.f);
    return 26;
  } catch (std::runtime_error const &) {
  }
  return 0;
}

int host_side_semantics() {
  if (int rc = random_initialisers()) return rc;
  {  // the generated value types (eq, ord, hash)
    std::unordered_set<legion_dim_t> dims{legion_dim_t{1}, legion_dim_t{1}, legion_dim_t{2}};
    std::unordered_set<ProfilingSettings> ps{ProfilingSettings{1, 2}, ProfilingSettings{1, 2}};
    if (dims.size() != 2 || ps.size() != 1 || !(legion_dim_t{1} < legion_dim_t{2}) ||
        !(ProfilingSettings{1, 2} < ProfilingSettings{1, 3}) || ProfilingSettings{0, 1} != ProfilingSettings{0, 1})
      return 30;
  }
  if (DataTypeDispatch1<SizeOf>{}(DataType::DOUBLE, size_t(3)) != 24 || dispatch<SizeOf>(DataType::INT32, size_t(2)) != 8) return 10;
  int widen = 0;
  DataTypeDispatch2<Widen>{}(DataType::FLOAT, DataType::INT64, &widen);
  if (widen != 4) return 11;
  try {
    DataTypeDispatch1<SizeOf>{}(DataType::HALF, size_t(1));   // not a dispatchable type in the reference either
    return 12;
  } catch (std::runtime_error const &) {
  }
  {
    double d[2] = {1.0, 2.0};
    GenericTensorAccessorW wd{{DataType::DOUBLE, ArrayShape(std::vector<size_t>{2}), d}};
    if (get<DataType::DOUBLE>(wd) != d || get_double_ptrs(std::vector<GenericTensorAccessorW>{wd, wd}).size() != 2) return 13;
    if (get_shape_and_datatype(wd).second != DataType::DOUBLE) return 14;
    try {
      get_float_ptr(wd);
      return 15;
    } catch (std::runtime_error const &) {
    }
  }
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
