// Host-side runtime of the kernel API: slab/size-class device arena, per-device handle,
// streams, event timers, fills.
#include "common.cuh"

#include <algorithm>
#include <map>
#include <unordered_map>
#include <vector>

struct ffk_arena {
  size_t slab_bytes;
  struct Slab {
    char *base;
    size_t size, used;
  };
  std::vector<Slab> slabs;
  std::vector<void *> big;                                  // direct driver allocations
  std::map<size_t, std::vector<void *>> free_lists;         // size class -> recycled blocks
  std::unordered_map<void *, size_t> live;                  // block -> size class
  size_t in_use = 0, reserved = 0, driver_allocs = 0;
};

// struct ffk_timer lives in common.cuh (the GEMM launcher records into it for ffk_gemm_time_next)

namespace ffk {
namespace {

size_t size_class(size_t bytes) {
  size_t c = 256;
  while (c < bytes) c <<= 1;
  return c;
}

template <typename T>
__global__ void __launch_bounds__(256) fill_value_kernel(T *p, size_t n, T v) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_arena_create(ffk_arena_t **out, size_t slab_bytes) {
  FFK_REQUIRE(out, "arena_create: null out pointer");
  auto *a = new ffk_arena();
  a->slab_bytes = slab_bytes ? std::max<size_t>(slab_bytes, 1 << 20) : (size_t(256) << 20);
  *out = a;
  return FFK_OK;
}

int ffk_arena_allocate(ffk_arena_t *a, size_t bytes, void **out) {
  FFK_REQUIRE(a && out, "arena_allocate: null argument");
  if (bytes == 0) bytes = 1;
  size_t cls = size_class(bytes);
  auto &fl = a->free_lists[cls];
  void *p = nullptr;
  if (!fl.empty()) {
    p = fl.back();
    fl.pop_back();
  } else if (cls > a->slab_bytes / 2) {
    FFK_CUDA(cudaMalloc(&p, cls));
    a->big.push_back(p);
    a->reserved += cls;
    a->driver_allocs += 1;
  } else {
    if (a->slabs.empty() || a->slabs.back().used + cls > a->slabs.back().size) {
      char *base = nullptr;
      FFK_CUDA(cudaMalloc(&base, a->slab_bytes));
      a->slabs.push_back({base, a->slab_bytes, 0});
      a->reserved += a->slab_bytes;
      a->driver_allocs += 1;
    }
    auto &s = a->slabs.back();
    p = s.base + s.used;
    s.used += cls;
  }
  a->live[p] = cls;
  a->in_use += cls;
  *out = p;
  return FFK_OK;
}

int ffk_arena_deallocate(ffk_arena_t *a, void *ptr) {
  FFK_REQUIRE(a, "arena_deallocate: null arena");
  if (!ptr) return FFK_OK;
  auto it = a->live.find(ptr);
  FFK_REQUIRE(it != a->live.end(), "arena_deallocate: pointer %p was not allocated here", ptr);
  size_t cls = it->second;  // read the size BEFORE erasing (the reference's TrackedAllocator
                            // subtracts an uninitialised size: tracked_allocator.cc:15-20)
  a->live.erase(it);
  a->in_use -= cls;
  a->free_lists[cls].push_back(ptr);
  return FFK_OK;
}

int ffk_arena_destroy(ffk_arena_t *a) {
  if (!a) return FFK_OK;
  for (auto &s : a->slabs) cudaFree(s.base);
  for (void *p : a->big) cudaFree(p);
  delete a;
  return FFK_OK;
}

size_t ffk_arena_bytes_in_use(ffk_arena_t const *a) { return a ? a->in_use : 0; }
size_t ffk_arena_bytes_reserved(ffk_arena_t const *a) { return a ? a->reserved : 0; }
size_t ffk_arena_driver_allocations(ffk_arena_t const *a) { return a ? a->driver_allocs : 0; }

int ffk_handle_create(ffk_handle_t *out, size_t workspace_bytes, int allow_conv) {
  FFK_REQUIRE(out, "handle_create: null out pointer");
  std::memset(out, 0, sizeof(*out));
  out->workSpaceSize = workspace_bytes ? workspace_bytes : size_t(1024) * 1024;
  out->allowTensorOpMathConversion = allow_conv;
  FFK_CUDA(cudaMalloc(&out->workSpace, out->workSpaceSize));
  return FFK_OK;
}

int ffk_handle_destroy(ffk_handle_t *h) {
  if (!h) return FFK_OK;
  if (h->workSpace) FFK_CUDA(cudaFree(h->workSpace));
  h->workSpace = nullptr;
  h->workSpaceSize = 0;
  return FFK_OK;
}

int ffk_stream_create(void **out_stream) {
  FFK_REQUIRE(out_stream, "stream_create: null out pointer");
  cudaStream_t s;
  FFK_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out_stream = s;
  return FFK_OK;
}
int ffk_stream_destroy(void *stream) {
  if (stream) FFK_CUDA(cudaStreamDestroy(as_stream(stream)));
  return FFK_OK;
}
int ffk_stream_synchronize(void *stream) {
  FFK_CUDA(cudaStreamSynchronize(as_stream(stream)));
  return FFK_OK;
}

int ffk_timer_create(ffk_timer_t **out) {
  FFK_REQUIRE(out, "timer_create: null out pointer");
  auto *t = new ffk_timer();
  cudaError_t e = cudaEventCreate(&t->start);
  if (e == cudaSuccess) e = cudaEventCreate(&t->stop);
  if (e != cudaSuccess) {
    delete t;
    return cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__);
  }
  *out = t;
  return FFK_OK;
}
int ffk_timer_begin(ffk_timer_t *t, void *stream) {
  FFK_REQUIRE(t, "timer_begin: null timer");
  FFK_CUDA(cudaEventRecord(t->start, as_stream(stream)));
  return FFK_OK;
}
int ffk_timer_end(ffk_timer_t *t, void *stream, float *elapsed_ms) {
  FFK_REQUIRE(t && elapsed_ms, "timer_end: null argument");
  FFK_CUDA(cudaEventRecord(t->stop, as_stream(stream)));
  FFK_CUDA(cudaEventSynchronize(t->stop));
  FFK_CUDA(cudaEventElapsedTime(elapsed_ms, t->start, t->stop));
  return FFK_OK;
}
int ffk_timer_read(ffk_timer_t *t, float *elapsed_ms) {
  FFK_REQUIRE(t && elapsed_ms, "timer_read: null argument");
  FFK_CUDA(cudaEventSynchronize(t->stop));
  FFK_CUDA(cudaEventElapsedTime(elapsed_ms, t->start, t->stop));
  return FFK_OK;
}
int ffk_gemm_time_next(ffk_timer_t *t) {
  ffk::gemm_timer_slot() = t;
  return FFK_OK;
}
int ffk_timer_destroy(ffk_timer_t *t) {
  if (!t) return FFK_OK;
  cudaEventDestroy(t->start);
  cudaEventDestroy(t->stop);
  delete t;
  return FFK_OK;
}

int ffk_fill(void *stream, void *ptr, size_t n, double value, int dtype) {
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(ptr, "fill: null pointer");
  cudaStream_t s = as_stream(stream);
  unsigned grid = stream_grid(n, 256, 8);
  switch (dtype) {
    case FFK_BOOL: fill_value_kernel<bool><<<grid, 256, 0, s>>>((bool *)ptr, n, value != 0.0); break;
    case FFK_INT32: fill_value_kernel<int32_t><<<grid, 256, 0, s>>>((int32_t *)ptr, n, (int32_t)value); break;
    case FFK_INT64: fill_value_kernel<int64_t><<<grid, 256, 0, s>>>((int64_t *)ptr, n, (int64_t)value); break;
    case FFK_HALF: fill_value_kernel<__half><<<grid, 256, 0, s>>>((__half *)ptr, n, __float2half((float)value)); break;
    case FFK_FLOAT: fill_value_kernel<float><<<grid, 256, 0, s>>>((float *)ptr, n, (float)value); break;
    case FFK_DOUBLE: fill_value_kernel<double><<<grid, 256, 0, s>>>((double *)ptr, n, value); break;
    case FFK_BFLOAT16: fill_value_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((__nv_bfloat16 *)ptr, n, __float2bfloat16((float)value)); break;
    default: set_error("fill: bad dtype %d", dtype); return FFK_ERR_DTYPE;
  }
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // extern "C"
