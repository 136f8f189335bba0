// C++ veneer, part 2: allocation, handle, stream and profiling objects of the kernel API
// (allocation.h, local_cuda_allocator.h, managed_ff_stream.h, managed_per_device_ff_handle.h,
// profiling.h of the reference) over the ffk runtime C ABI.
#pragma once
#include "types.h"

namespace FlexFlow {

// ---- allocation.h:10-35
struct IAllocator {
  virtual void *allocate(size_t) = 0;
  virtual void deallocate(void *) = 0;
  virtual ~IAllocator() = default;
};

struct Allocator {
  Allocator() = delete;
  Allocator(std::shared_ptr<IAllocator> ptr) : i_allocator(std::move(ptr)) {}
  GenericTensorAccessorW allocate_tensor(TensorShape const &shape) {
    void *p = this->allocate(get_size_in_bytes(shape));
    return GenericTensorAccessorW{{shape.data_type, ArrayShape(shape), p}};
  }
  void *allocate(size_t mem_size) { return i_allocator->allocate(mem_size); }
  void deallocate(void *ptr) { i_allocator->deallocate(ptr); }
  template <typename T, typename... Args>
  static std::enable_if_t<std::is_base_of<IAllocator, T>::value, Allocator> create(Args &&...args) {
    return Allocator(std::make_shared<T>(std::forward<Args>(args)...));
  }

private:
  std::shared_ptr<IAllocator> i_allocator;
};

// local_cuda_allocator.h:6-20. Same interface; blocks come from the slab arena in libffk.so
// instead of one cudaMalloc per tensor, and deallocate() recycles them.
struct LocalCudaAllocator : public IAllocator {
  LocalCudaAllocator() { FFK_CHECK(ffk_arena_create(&arena, 0)); }
  LocalCudaAllocator(LocalCudaAllocator const &) = delete;
  LocalCudaAllocator(LocalCudaAllocator &&) = delete;
  ~LocalCudaAllocator() override { ffk_arena_destroy(arena); }
  void *allocate(size_t n) override {
    void *p = nullptr;
    FFK_CHECK(ffk_arena_allocate(arena, n, &p));
    return p;
  }
  void deallocate(void *p) override { FFK_CHECK(ffk_arena_deallocate(arena, p)); }
  size_t bytes_in_use() const { return ffk_arena_bytes_in_use(arena); }
  size_t driver_allocations() const { return ffk_arena_driver_allocations(arena); }

private:
  ffk_arena_t *arena = nullptr;
};
inline Allocator create_local_cuda_memory_allocator() { return Allocator::create<LocalCudaAllocator>(); }

// lib/local-execution/include/local-execution/tracked_allocator.h — counts live bytes for the
// cost estimator. (The reference's deallocate subtracts an uninitialised size,
// src/tracked_allocator.cc:15-20; here the size is looked up before the entry is erased.)
struct TrackedAllocator : public IAllocator {
  explicit TrackedAllocator(Allocator a) : inner(std::move(a)) {}
  void *allocate(size_t n) override {
    void *p = inner.allocate(n);
    sizes.emplace_back(p, n);
    current += n;
    return p;
  }
  void deallocate(void *p) override {
    for (size_t i = 0; i < sizes.size(); ++i)
      if (sizes[i].first == p) {
        current -= sizes[i].second;
        sizes.erase(sizes.begin() + i);
        break;
      }
    inner.deallocate(p);
  }
  size_t get_current_mem_usage() const { return current; }

private:
  Allocator inner;
  std::vector<std::pair<void *, size_t>> sizes;
  size_t current = 0;
};

// managed_ff_stream.h:8-25
struct ManagedFFStream {
  ManagedFFStream() : stream(new ffStream_t) {
    void *s = nullptr;
    FFK_CHECK(ffk_stream_create(&s));
    *stream = static_cast<ffStream_t>(s);
  }
  ManagedFFStream(ManagedFFStream const &) = delete;
  ManagedFFStream &operator=(ManagedFFStream const &) = delete;
  ManagedFFStream(ManagedFFStream &&o) noexcept : stream(std::exchange(o.stream, nullptr)) {}
  ManagedFFStream &operator=(ManagedFFStream &&o) noexcept {
    std::swap(stream, o.stream);
    return *this;
  }
  ~ManagedFFStream() {
    if (stream) {
      ffk_stream_destroy(*stream);
      delete stream;
    }
  }
  ffStream_t const &raw_stream() const { return *stream; }

private:
  ffStream_t *stream;
};

// managed_per_device_ff_handle.h:8-27 (1 MiB workspace, tensor-op conversion allowed).
// Behind the op workspace the handle also reserves GEMM scratch (256 MiB by default: 0.14 % of
// the GPU's memory) and registers it for this device with ffk_set_gemm_workspace; and because
// the reference's handle says allowTensorOpMathConversion = true (managed_per_device_ff_handle.cc:9)
// it switches ffk_set_gemm_fp32_tensor_math on. An UNCHANGED fp32 caller of Linear / BatchMatmul /
// MultiHeadAttention / Conv2D therefore runs its large products on the tensor cores (bf16 split,
// fp32-class accuracy). The scratch is kept apart from `workSpace` because an op's own scratch
// (e.g. Conv2D's im2col panel) can be a GEMM operand. gemm_scratch_bytes = 0 leaves fp32 products
// on the exact CUDA-core kernel.
struct ManagedPerDeviceFFHandle {
  static constexpr size_t kDefaultGemmScratch = size_t(256) << 20;
  explicit ManagedPerDeviceFFHandle(size_t workspace_bytes = 1024 * 1024,
                                    size_t gemm_scratch_bytes = kDefaultGemmScratch)
      : handle(new PerDeviceFFHandle) {
    ffk_handle_t h;
    size_t ws = (workspace_bytes + 255) / 256 * 256;
    FFK_CHECK(ffk_handle_create(&h, ws + gemm_scratch_bytes, 1));
    *handle = PerDeviceFFHandle{nullptr, nullptr, h.workSpace, ws, true, nullptr};
    if (gemm_scratch_bytes) {
      FFK_CHECK(ffk_set_gemm_workspace(static_cast<char *>(h.workSpace) + ws, gemm_scratch_bytes));
      FFK_CHECK(ffk_set_gemm_fp32_tensor_math(handle->allowTensorOpMathConversion ? 1 : 0));
      owns_gemm_scratch = true;
    }
  }
  ManagedPerDeviceFFHandle(ManagedPerDeviceFFHandle const &) = delete;
  ManagedPerDeviceFFHandle &operator=(ManagedPerDeviceFFHandle const &) = delete;
  ManagedPerDeviceFFHandle(ManagedPerDeviceFFHandle &&o) noexcept
      : handle(std::exchange(o.handle, nullptr)), owns_gemm_scratch(std::exchange(o.owns_gemm_scratch, false)) {}
  ManagedPerDeviceFFHandle &operator=(ManagedPerDeviceFFHandle &&o) noexcept {
    std::swap(handle, o.handle);
    std::swap(owns_gemm_scratch, o.owns_gemm_scratch);
    return *this;
  }
  ~ManagedPerDeviceFFHandle() {
    if (handle) {
      if (owns_gemm_scratch) {
        ffk_set_gemm_workspace(nullptr, 0);
        ffk_set_gemm_fp32_tensor_math(0);
      }
      ffk_handle_t h{nullptr, nullptr, handle->workSpace, handle->workSpaceSize, 1, nullptr};
      ffk_handle_destroy(&h);
      delete handle;
    }
  }
  PerDeviceFFHandle const &raw_handle() const { return *handle; }

private:
  PerDeviceFFHandle *handle;
  bool owns_gemm_scratch = false;
};

// The stream every profiled call runs on. The reference's get_legion_stream creates a new
// stream on every call and never destroys it (cuda_helper.cu:8-14); here one stream per host
// thread is created on first use and reused.
inline cudaError_t get_legion_stream(cudaStream_t *stream) {
  thread_local ManagedFFStream s;
  *stream = s.raw_stream();
  return cudaSuccess;
}

// profiling.h:24-49: run f(stream, args...) warmup+measure times and return the elapsed
// milliseconds of the MEASURED iterations in total (not their mean), which is the number the
// local cost estimator sums up.
template <typename F, typename... Ts>
std::optional<float> profiling_wrapper(F const &f, ProfilingSettings const &settings, Ts &&...ts) {
  ffStream_t stream;
  checkCUDA(get_legion_stream(&stream));
  thread_local struct Timer {
    ffk_timer_t *t = nullptr;
    Timer() { FFK_CHECK(ffk_timer_create(&t)); }
    ~Timer() { ffk_timer_destroy(t); }
  } timer;
  for (int i = 0; i < settings.warmup_iters + settings.measure_iters; i++) {
    if (i == settings.warmup_iters) FFK_CHECK(ffk_timer_begin(timer.t, stream));
    f(stream, std::forward<Ts>(ts)...);
  }
  float elapsed = 0;
  if (settings.measure_iters <= 0) FFK_CHECK(ffk_timer_begin(timer.t, stream));
  FFK_CHECK(ffk_timer_end(timer.t, stream, &elapsed));
  return elapsed;
}
template <typename F, typename... Ts>
std::optional<float> profiling_wrapper(F const &f, bool enable_profiling, Ts &&...ts) {
  if (enable_profiling) {
    ProfilingSettings settings = {0, 1};
    return profiling_wrapper<F, Ts...>(f, settings, std::forward<Ts>(ts)...);
  }
  ffStream_t stream;
  checkCUDA(get_legion_stream(&stream));
  f(stream, std::forward<Ts>(ts)...);
  return std::nullopt;
}

// perf_metrics.h:9-41 — host-side value type; the device accumulator is ffk_perf_metrics_t.
struct PerfMetrics {
  PerfMetrics() = delete;
  explicit PerfMetrics(double start_time_) : start_time(start_time_), current_time(start_time_) {}
  int train_all = 0;
  std::optional<int> train_correct = 0;
  std::optional<float> cce_loss = std::nullopt;
  std::optional<float> sparse_cce_loss = 0.0f;
  std::optional<float> mse_loss = 0.0f;
  std::optional<float> rmse_loss = 0.0f;
  std::optional<float> mae_loss = 0.0f;
  double start_time;
  double current_time;
};
inline float get_accuracy(PerfMetrics const &m) {
  return m.train_all ? 100.0f * float(m.train_correct.value_or(0)) / float(m.train_all) : 0.f;
}
inline float get_throughput(PerfMetrics const &m) {
  double dt = m.current_time - m.start_time;
  return dt > 0 ? float(m.train_all / (dt * 1e-6)) : 0.f;  // times are in microseconds
}

}  // namespace FlexFlow
