// Loss gradients, optimizer updates, metric accumulation. All single-pass streaming kernels.
#include "common.cuh"

namespace ffk {
namespace {

constexpr int kThreads = 256;

// g = (p - onehot(label)) * s over the flat [rows][classes] tensor in 16-byte packets (a packet may
// straddle rows: one division per packet finds the first element's row, the rest is stepped).
template <typename T>
__global__ void __launch_bounds__(kThreads)
    sparse_cce_bwd_kernel(T *__restrict__ g, T const *__restrict__ p, int const *__restrict__ label,
                          long long rows, int classes, int k, float s, bool vec_ok) {
  constexpr int VN = Vec16<T>::N;
  long long const n = rows * classes, stride = (long long)gridDim.x * blockDim.x;
  long long const tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long const nvec = vec_ok ? n / VN : 0;
  for (long long i = tid; i < nvec; i += stride) {
    long long const e0 = i * VN;
    long long r = e0 / classes;
    int c = int(e0 - r * classes);
    int lab = label[r / k];
    Vec16<T> a = ld_stream(p + e0), o;
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      float v = to_f32(a.v[j]);
      if (c == lab) v -= 1.0f;
      o.v[j] = from_f32<T>(v * s);
      if (++c == classes) {
        c = 0;
        ++r;
        if (r < rows) lab = label[r / k];
      }
    }
    st_stream(g + e0, o);
  }
  for (long long i = nvec * VN + tid; i < n; i += stride) {
    long long r = i / classes;
    int c = int(i - r * classes);
    float v = to_f32(p[i]);
    if (c == label[r / k]) v -= 1.0f;
    g[i] = from_f32<T>(v * s);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    diff_scale_kernel(T *__restrict__ g, T const *__restrict__ p, T const *__restrict__ y, size_t n,
                      float s, bool vec_ok) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = vec_ok ? n / VN : 0, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> a = ld_stream(p + i * VN), b = ld_stream(y + i * VN), o;
#pragma unroll
    for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>((to_f32(a.v[j]) - to_f32(b.v[j])) * s);
    st_stream(g + i * VN, o);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride)
    g[i] = from_f32<T>((to_f32(p[i]) - to_f32(y[i])) * s);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) fill_kernel(T *__restrict__ g, size_t n, float v) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    g[i] = from_f32<T>(v);
}

// ------------------------------------------------------------------ optimizers
template <typename T>
__device__ __forceinline__ float replica_sum(T const *__restrict__ g, size_t i, size_t size,
                                             int reps) {
  float acc = to_f32(g[i]);
  for (int r = 1; r < reps; ++r) acc += to_f32(g[size_t(r) * size + i]);
  return acc;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    sgd_kernel(size_t n, float lr, float wd, float mom, bool nesterov, T const *__restrict__ g,
               int reps, float *__restrict__ V, T *__restrict__ W, float *__restrict__ master) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float w = master ? master[i] : to_f32(W[i]);
    float gt = replica_sum(g, i, n, reps) + wd * w;
    if (mom > 0.0f) {
      float v = V[i] * mom + gt;
      V[i] = v;
      gt = nesterov ? gt + mom * v : v;
    }
    w -= lr * gt;
    if (master) master[i] = w;
    W[i] = from_f32<T>(w);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    adam_kernel(size_t n, float alpha_t, float b1, float b2, float wd, float eps,
                T const *__restrict__ g, int reps, float *__restrict__ M, float *__restrict__ V,
                T *__restrict__ W, float *__restrict__ master) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float w = master ? master[i] : to_f32(W[i]);
    float gt = replica_sum(g, i, n, reps) + wd * w;
    float mt = b1 * M[i] + (1 - b1) * gt;
    float vt = b2 * V[i] + (1 - b2) * gt * gt;
    M[i] = mt;
    V[i] = vt;
    w -= alpha_t * mt / (sqrtf(vt) + eps);
    if (master) master[i] = w;
    W[i] = from_f32<T>(w);
  }
}

// ------------------------------------------------------------------ metrics
constexpr float kLogMin = 0.00000001f;

struct Partial {
  int all, correct;
  float cce, scce, mse, rmse, mae;
};

__device__ __forceinline__ void block_commit(Partial p, ffk_perf_metrics_t *perf) {
  // warp-shuffle then shared-memory reduction: one set of atomics per CTA instead of per sample
  __shared__ Partial sh[kThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    p.all += __shfl_xor_sync(0xffffffffu, p.all, o);
    p.correct += __shfl_xor_sync(0xffffffffu, p.correct, o);
    p.cce += __shfl_xor_sync(0xffffffffu, p.cce, o);
    p.scce += __shfl_xor_sync(0xffffffffu, p.scce, o);
    p.mse += __shfl_xor_sync(0xffffffffu, p.mse, o);
    p.rmse += __shfl_xor_sync(0xffffffffu, p.rmse, o);
    p.mae += __shfl_xor_sync(0xffffffffu, p.mae, o);
  }
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = p;
  __syncthreads();
  if (threadIdx.x == 0) {
    Partial t = sh[0];
    for (int w = 1; w < kThreads / 32; ++w) {
      t.all += sh[w].all; t.correct += sh[w].correct; t.cce += sh[w].cce; t.scce += sh[w].scce;
      t.mse += sh[w].mse; t.rmse += sh[w].rmse; t.mae += sh[w].mae;
    }
    if (t.all) atomicAdd(&perf->train_all, t.all);
    if (t.correct) atomicAdd(&perf->train_correct, t.correct);
    if (t.cce != 0.f) atomicAdd(&perf->cce_loss, t.cce);
    if (t.scce != 0.f) atomicAdd(&perf->sparse_cce_loss, t.scce);
    if (t.mse != 0.f) atomicAdd(&perf->mse_loss, t.mse);
    if (t.rmse != 0.f) atomicAdd(&perf->rmse_loss, t.rmse);
    if (t.mae != 0.f) atomicAdd(&perf->mae_loss, t.mae);
  }
}

__global__ void __launch_bounds__(kThreads)
    metrics_sparse_kernel(float const *__restrict__ logits, int const *__restrict__ labels,
                          ffk_perf_metrics_t *perf, int metrics, int samples, int classes) {
  Partial p{};
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < samples; b += gridDim.x * blockDim.x) {
    float const *row = logits + size_t(b) * classes;
    if (metrics & FFK_METRIC_ACCURACY) {
      float mx = -1.0f;
      int arg = -1;
      for (int i = 0; i < classes; ++i)
        if (row[i] > mx) { mx = row[i]; arg = i; }
      p.all += 1;
      if (labels[b] == arg) p.correct += 1;
    }
    if (metrics & FFK_METRIC_SPARSE_CCE) p.scce += -logf(fmaxf(row[labels[b]], kLogMin));
    if (metrics & (FFK_METRIC_MSE | FFK_METRIC_RMSE | FFK_METRIC_MAE)) {
      float mse = 0.f, mae = 0.f;
      for (int i = 0; i < classes; ++i) {
        float d = row[i] - (labels[b] == i ? 1.0f : 0.0f);
        mse += d * d;
        mae += fabsf(d);
      }
      if (metrics & FFK_METRIC_MSE) p.mse += mse;
      if (metrics & FFK_METRIC_RMSE) p.rmse += sqrtf(mse);
      if (metrics & FFK_METRIC_MAE) p.mae += mae;
    }
  }
  block_commit(p, perf);
}

__global__ void __launch_bounds__(kThreads)
    metrics_label_kernel(float const *__restrict__ logits, float const *__restrict__ labels,
                         ffk_perf_metrics_t *perf, int metrics, int samples, int classes) {
  Partial p{};
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < samples; b += gridDim.x * blockDim.x) {
    float const *row = logits + size_t(b) * classes;
    float const *lab = labels + size_t(b) * classes;
    p.all += 1;
    if (metrics & FFK_METRIC_ACCURACY) {
      if (classes == 1) {
        p.all += 1;
        p.correct += 1;
      } else {
        float mx = 0.f;
        int arg = -1, truth = -1;
        for (int i = 0; i < classes; ++i) {
          if (arg == -1 || row[i] > mx) { mx = row[i]; arg = i; }
          if (lab[i] > 0.9f) truth = i;
        }
        if (truth == arg) p.correct += 1;
      }
    }
    if (metrics & FFK_METRIC_CCE) {
      float cce = 0.f;
      for (int i = 0; i < classes; ++i)
        if (lab[i] > 0.f) cce += lab[i] * -logf(fmaxf(row[i], kLogMin));
      p.cce += cce;
    }
    if (metrics & (FFK_METRIC_MSE | FFK_METRIC_RMSE | FFK_METRIC_MAE)) {
      float mse = 0.f, mae = 0.f;
      for (int i = 0; i < classes; ++i) {
        float d = row[i] - lab[i];
        mse += d * d;
        mae += fabsf(d);
      }
      if (metrics & FFK_METRIC_MSE) p.mse += mse;
      if (metrics & FFK_METRIC_RMSE) p.rmse += sqrtf(mse);
      if (metrics & FFK_METRIC_MAE) p.mae += mae;
    }
  }
  block_commit(p, perf);
}

// Wide rows (>= 64 classes): one WARP per sample so that the row is read with coalesced
// loads; per-lane partial results are combined with shuffles. Same definitions as the
// thread-per-sample kernels above (first maximum wins, last label above 0.9 is the truth).
__device__ __forceinline__ void warp_argmax(float &mx, int &arg) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, mx, o);
    int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    bool take = oa >= 0 && (arg < 0 || om > mx || (om == mx && oa < arg));
    if (take) { mx = om; arg = oa; }
  }
}

template <bool SPARSE>
__global__ void __launch_bounds__(kThreads)
    metrics_wide_kernel(float const *__restrict__ logits, void const *__restrict__ labels_,
                        ffk_perf_metrics_t *perf, int metrics, int samples, int classes) {
  Partial p{};
  int const lane = threadIdx.x & 31;
  int const warps = gridDim.x * (kThreads / 32);
  for (int b = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); b < samples; b += warps) {
    float const *row = logits + size_t(b) * classes;
    float const *lab = SPARSE ? nullptr : reinterpret_cast<float const *>(labels_) + size_t(b) * classes;
    int const label = SPARSE ? reinterpret_cast<int const *>(labels_)[b] : -1;
    float mx = SPARSE ? -1.0f : 0.f, cce = 0.f, mse = 0.f, mae = 0.f;
    int arg = -1, truth = -1;
    for (int i = lane; i < classes; i += 32) {
      float const v = row[i];
      float const t = SPARSE ? (label == i ? 1.0f : 0.0f) : lab[i];
      if (SPARSE ? (v > mx) : (arg == -1 || v > mx)) { mx = v; arg = i; }
      if (!SPARSE && t > 0.9f) truth = i;
      if (!SPARSE && t > 0.f) cce += t * -logf(fmaxf(v, kLogMin));
      float const d = v - t;
      mse += d * d;
      mae += fabsf(d);
    }
    warp_argmax(mx, arg);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) truth = max(truth, __shfl_xor_sync(0xffffffffu, truth, o));
    cce = warp_sum(cce);
    mse = warp_sum(mse);
    mae = warp_sum(mae);
    if (lane == 0) {
      if (SPARSE) {
        if (metrics & FFK_METRIC_ACCURACY) {
          p.all += 1;
          if (label == arg) p.correct += 1;
        }
        if (metrics & FFK_METRIC_SPARSE_CCE) p.scce += -logf(fmaxf(row[label], kLogMin));
      } else {
        p.all += 1;
        if ((metrics & FFK_METRIC_ACCURACY) && truth == arg) p.correct += 1;
        if (metrics & FFK_METRIC_CCE) p.cce += cce;
      }
      if (metrics & FFK_METRIC_MSE) p.mse += mse;
      if (metrics & FFK_METRIC_RMSE) p.rmse += sqrtf(mse);
      if (metrics & FFK_METRIC_MAE) p.mae += mae;
    }
  }
  block_commit(p, perf);
}

constexpr int kWideMetricsClasses = 64;

int diff_scale(void *stream, void *g, void const *p, void const *y, size_t n, float s, int dtype) {
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(g && p && y, "loss backward: null tensor");
  FFK_DISPATCH_FLOAT(dtype, T, {
    bool vec_ok = aligned16(g) && aligned16(p) && aligned16(y);
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, 8);
    diff_scale_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>((T *)g, (T const *)p,
                                                                   (T const *)y, n, s, vec_ok);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_sparse_categorical_crossentropy_loss_backward(void *stream, void *logit_grad,
                                                      void const *logit, int const *label,
                                                      int num_samples, int num_classes, int k,
                                                      float scale_factor, int dtype) {
  FFK_REQUIRE(num_samples >= 0 && num_classes > 0 && k > 0, "sparse cce: bad extents");
  if (num_samples == 0) return FFK_OK;
  FFK_REQUIRE(logit_grad && logit && label, "sparse cce: null tensor");
  FFK_DISPATCH_FLOAT(dtype, T, {
    bool vec_ok = aligned16(logit_grad) && aligned16(logit);
    unsigned grid = stream_grid(size_t(num_samples) * num_classes / Vec16<T>::N + 1, kThreads, 8);
    sparse_cce_bwd_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        (T *)logit_grad, (T const *)logit, label, num_samples, num_classes, k,
        scale_factor * float(k), vec_ok);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_categorical_crossentropy_loss_backward(void *stream, void *logit_grad, void const *logit,
                                               void const *label, size_t volume,
                                               float scale_factor, int dtype) {
  return diff_scale(stream, logit_grad, logit, label, volume, scale_factor, dtype);
}

int ffk_mean_squared_error_avg_loss_backward(void *stream, void *logit_grad, void const *logit,
                                             void const *label, size_t volume,
                                             float scale_factor, int dtype) {
  return diff_scale(stream, logit_grad, logit, label, volume, scale_factor, dtype);
}

int ffk_identity_loss_backward(void *stream, void *loss_grad, void const *loss, size_t volume,
                               float scale_factor, int dtype) {
  (void)loss;
  if (volume == 0) return FFK_OK;
  FFK_REQUIRE(loss_grad, "identity loss: null tensor");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(volume, kThreads, 8);
    fill_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>((T *)loss_grad, volume,
                                                             1.0f * scale_factor);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_sgd_update(void *stream, float lr, float momentum, int nesterov, float weight_decay,
                   void const *weight_grad, size_t size, int num_replicas, void *weight,
                   float *sgd_v, int dtype, float *master_weight) {
  FFK_REQUIRE(num_replicas >= 1, "sgd: num_replicas must be >= 1");
  if (size == 0) return FFK_OK;
  FFK_REQUIRE(weight_grad && weight, "sgd: null tensor");
  FFK_REQUIRE(momentum <= 0.f || sgd_v, "sgd: momentum needs the velocity buffer");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(size, kThreads, 8);
    sgd_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        size, lr, weight_decay, momentum, nesterov != 0, (T const *)weight_grad, num_replicas,
        sgd_v, (T *)weight, master_weight);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_adam_update(void *stream, float alpha_t, float beta1, float beta2, float weight_decay,
                    float epsilon, void const *weight_grad, size_t size, int num_replicas,
                    float *adam_m, float *adam_v, void *weight, int dtype, float *master_weight) {
  FFK_REQUIRE(num_replicas >= 1, "adam: num_replicas must be >= 1");
  if (size == 0) return FFK_OK;
  FFK_REQUIRE(weight_grad && weight && adam_m && adam_v, "adam: null tensor");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(size, kThreads, 8);
    adam_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        size, alpha_t, beta1, beta2, weight_decay, epsilon, (T const *)weight_grad, num_replicas,
        adam_m, adam_v, (T *)weight, master_weight);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_update_metrics_sparse_label(void *stream, int metrics, float const *logit,
                                    int const *label, int num_samples, int num_classes,
                                    ffk_perf_metrics_t *perf) {
  FFK_REQUIRE(num_samples >= 0 && num_classes > 0, "metrics: bad extents");
  if (num_samples == 0) return FFK_OK;
  FFK_REQUIRE(logit && label && perf, "metrics: null pointer");
  if (num_classes >= kWideMetricsClasses) {
    unsigned grid = stream_grid(size_t(num_samples) * 32, kThreads, 8);
    metrics_wide_kernel<true><<<grid, kThreads, 0, as_stream(stream)>>>(logit, label, perf, metrics,
                                                                        num_samples, num_classes);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  unsigned grid = stream_grid(num_samples, kThreads, 4);
  metrics_sparse_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(logit, label, perf, metrics,
                                                                  num_samples, num_classes);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_update_metrics_label(void *stream, int metrics, float const *logit, float const *label,
                             int num_samples, int num_classes, ffk_perf_metrics_t *perf) {
  FFK_REQUIRE(num_samples >= 0 && num_classes > 0, "metrics: bad extents");
  if (num_samples == 0) return FFK_OK;
  FFK_REQUIRE(logit && label && perf, "metrics: null pointer");
  if (num_classes >= kWideMetricsClasses) {
    unsigned grid = stream_grid(size_t(num_samples) * 32, kThreads, 8);
    metrics_wide_kernel<false><<<grid, kThreads, 0, as_stream(stream)>>>(logit, label, perf, metrics,
                                                                         num_samples, num_classes);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  unsigned grid = stream_grid(num_samples, kThreads, 4);
  metrics_label_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(logit, label, perf, metrics,
                                                                 num_samples, num_classes);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // extern "C"
