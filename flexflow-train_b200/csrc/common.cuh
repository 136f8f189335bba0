// Shared device/host helpers for the sm_100a kernel library (libffk.so).
// Everything here is ours; nothing is taken from the reference tree.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/ffk.h"

// Event pair of ffk_timer_* (include/ffk_runtime.h).
struct ffk_timer {
  cudaEvent_t start, stop;
};

namespace ffk {

// ---------------------------------------------------------------- errors
// C-ABI functions return 0 on success and a negative code on failure; the
// message is kept per host thread and read back with ffk_last_error().
void set_error(char const *fmt, ...);
int cuda_fail(cudaError_t e, char const *what, char const *file, int line);

#define FFK_CUDA(expr)                                                         \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess) {                                                   \
      return ::ffk::cuda_fail(_e, #expr, __FILE__, __LINE__);                  \
    }                                                                          \
  } while (0)

#define FFK_REQUIRE(cond, ...)                                                 \
  do {                                                                         \
    if (!(cond)) {                                                             \
      ::ffk::set_error(__VA_ARGS__);                                           \
      return FFK_ERR_INVALID;                                                  \
    }                                                                          \
  } while (0)

#define FFK_LAUNCH_CHECK() FFK_CUDA(cudaGetLastError())

// ffk_gemm_time_next: the timer (of the calling host thread) that the next tensor-core GEMM launch
// brackets with its start / stop events; cleared by that launch.
::ffk_timer *&gemm_timer_slot();

// ---------------------------------------------------------------- device
struct DeviceInfo {
  int sm_count;
  int max_smem_optin;
};
DeviceInfo const &device_info();

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kSMsB200 = 148;

// Grid for a streaming kernel: enough CTAs for `resident` CTAs on every SM, but
// never more than the work needs.
inline unsigned stream_grid(size_t work_items, int threads, int resident) {
  size_t need = (work_items + threads - 1) / threads;
  size_t cap = static_cast<size_t>(device_info().sm_count) * resident;
  if (need < 1) need = 1;
  return static_cast<unsigned>(need < cap ? need : cap);
}

inline size_t dtype_size(int dt) {
  switch (dt) {
    case FFK_BOOL: return 1;
    case FFK_INT32: return 4;
    case FFK_INT64: return 8;
    case FFK_HALF: return 2;
    case FFK_FLOAT: return 4;
    case FFK_DOUBLE: return 8;
    case FFK_BFLOAT16: return 2;
    default: return 0;
  }
}

// ---------------------------------------------------------------- numerics
template <typename T> __device__ __forceinline__ float to_f32(T v) { return static_cast<float>(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v) { return static_cast<T>(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// A 16-byte packet of T. All streaming kernels move data in these.
template <typename T> struct alignas(16) Vec16 {
  static constexpr int N = 16 / sizeof(T);
  T v[N];
};

template <typename T>
__device__ __forceinline__ Vec16<T> ld_stream(T const *p) {
  // read-once data: do not allocate in L1
  Vec16<T> r;
  int4 raw;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w)
               : "l"(p));
  *reinterpret_cast<int4 *>(&r) = raw;
  return r;
}
template <typename T>
__device__ __forceinline__ Vec16<T> ld_vec(T const *p) {
  Vec16<T> r;
  *reinterpret_cast<int4 *>(&r) = *reinterpret_cast<int4 const *>(p);
  return r;
}
template <typename T>
__device__ __forceinline__ void st_vec(T *p, Vec16<T> const &r) {
  *reinterpret_cast<int4 *>(p) = *reinterpret_cast<int4 const *>(&r);
}
template <typename T>
__device__ __forceinline__ void st_stream(T *p, Vec16<T> const &r) {
  int4 raw = *reinterpret_cast<int4 const *>(&r);
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(raw.x),
               "r"(raw.y), "r"(raw.z), "r"(raw.w)
               : "memory");
}

inline bool aligned16(void const *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024; `red` is >= 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float *red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float r = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (w == 0) r = warp_sum(r);
  if (threadIdx.x == 0) red[0] = r;
  __syncthreads();
  return red[0];
}
__device__ __forceinline__ float block_max(float v, float *red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float r = (threadIdx.x < nw) ? red[threadIdx.x] : -INFINITY;
  if (w == 0) r = warp_max(r);
  if (threadIdx.x == 0) red[0] = r;
  __syncthreads();
  return red[0];
}

// dtype dispatch for the floating kernels
#define FFK_DISPATCH_FLOAT(dt, T, ...)                                         \
  switch (dt) {                                                                \
    case FFK_FLOAT: { using T = float; __VA_ARGS__; break; }                   \
    case FFK_BFLOAT16: { using T = __nv_bfloat16; __VA_ARGS__; break; }        \
    case FFK_HALF: { using T = __half; __VA_ARGS__; break; }                   \
    default:                                                                   \
      ::ffk::set_error("%s: unsupported dtype %d", __func__, (int)(dt));       \
      return FFK_ERR_DTYPE;                                                    \
  }

}  // namespace ffk
