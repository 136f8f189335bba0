// Index-driven kernels: Embedding (row gather / bag sum, vectorised atomic scatter-add),
// Gather, TopK.
#include "common.cuh"

namespace ffk {
namespace {

constexpr int kThreads = 256;

// 16-byte reduction into global memory (sm_90+): one L2 atomic transaction per packet.
__device__ __forceinline__ void red_add_vec(float *p, Vec16<float> const &v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.v[0]), "f"(v.v[1]),
               "f"(v.v[2]), "f"(v.v[3])
               : "memory");
}
__device__ __forceinline__ void red_add_vec(__nv_bfloat16 *p, Vec16<__nv_bfloat16> const &v) {
  uint32_t const *w = reinterpret_cast<uint32_t const *>(&v);
  asm volatile("red.global.add.noftz.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]),
               "r"(w[1]), "r"(w[2]), "r"(w[3])
               : "memory");
}
__device__ __forceinline__ void red_add_vec(__half *p, Vec16<__half> const &v) {
  uint32_t const *w = reinterpret_cast<uint32_t const *>(&v);
  asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]),
               "r"(w[1]), "r"(w[2]), "r"(w[3])
               : "memory");
}
__device__ __forceinline__ void atomic_add_any(float *p, float v) { atomicAdd(p, v); }
__device__ __forceinline__ void atomic_add_any(double *p, double v) { atomicAdd(p, v); }
__device__ __forceinline__ void atomic_add_any(__half *p, __half v) { atomicAdd(p, v); }
__device__ __forceinline__ void atomic_add_any(__nv_bfloat16 *p, __nv_bfloat16 v) { atomicAdd(p, v); }

// ------------------------------------------------------------------ embedding forward
// One thread per (row, packet). Consecutive threads read consecutive 16-byte packets of
// one table row, so a 256-wide fp32 row is four fully coalesced 256-byte warp requests.
template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    embed_fwd_vec_kernel(TI const *__restrict__ idx, T *__restrict__ out,
                         T const *__restrict__ table, int ppr /*packets per row*/, int bag,
                         long long batch, float scale, bool aggregate) {
  constexpr int VN = Vec16<T>::N;
  long long const total = batch * ppr;
  long long const stride = (long long)gridDim.x * blockDim.x;
  long long const out_dim = (long long)ppr * VN;
  if (!aggregate) {
    // plain gather: four packets per thread and trip, a grid stride apart — the four index loads
    // go out together, then the four row loads, then the stores, so that every thread keeps four
    // independent 16-byte requests to the (randomly placed) table rows in flight
    constexpr int U = 4;
    bool const small = total < (1ll << 31);   // 32-bit divisions (64-bit ones run on the slow XU path)
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += stride * U) {
      long long b[U];
      int pk[U];
      long long row[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        long long const i = i0 + u * stride;
        b[u] = small ? (long long)(unsigned(i) / unsigned(ppr)) : i / ppr;
        pk[u] = int(i - b[u] * ppr);
        row[u] = i < total ? (long long)idx[b[u]] : 0;
      }
      Vec16<T> v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < total) v[u] = ld_vec(table + row[u] * out_dim + pk[u] * VN);  // rows may be re-read: keep in L2
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < total) st_stream(out + b[u] * out_dim + pk[u] * VN, v[u]);
    }
    return;
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long const b = i / ppr;
    int const pk = int(i - b * ppr);
    {
      float acc[VN];
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] = 0.f;
      for (int t = 0; t < bag; ++t) {
        long long const row = (long long)idx[b * bag + t];
        Vec16<T> v = ld_vec(table + row * out_dim + pk * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) acc[j] += to_f32(v.v[j]);
      }
      Vec16<T> o;
#pragma unroll
      for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>(acc[j] * scale);
      st_stream(out + b * out_dim + pk * VN, o);
    }
  }
}

template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    embed_fwd_scalar_kernel(TI const *__restrict__ idx, T *__restrict__ out,
                            T const *__restrict__ table, int out_dim, int bag, long long batch,
                            double scale, bool aggregate) {
  long long const total = batch * out_dim;
  long long const stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long const b = i / out_dim, off = i - b * out_dim;
    if (!aggregate) {
      out[i] = table[(long long)idx[b] * out_dim + off];
    } else {
      double acc = 0.0;
      for (int t = 0; t < bag; ++t)
        acc += (double)table[(long long)idx[b * bag + t] * out_dim + off];
      out[i] = (T)(acc * scale);
    }
  }
}

// ------------------------------------------------------------------ embedding backward
template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    embed_bwd_vec_kernel(TI const *__restrict__ idx, T const *__restrict__ dout,
                         T *__restrict__ dtable, int ppr, int bag, long long batch, float scale,
                         bool scaled) {
  constexpr int VN = Vec16<T>::N;
  long long const total = batch * ppr;
  long long const stride = (long long)gridDim.x * blockDim.x;
  long long const out_dim = (long long)ppr * VN;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long const b = i / ppr;
    int const pk = int(i - b * ppr);
    Vec16<T> g = ld_stream(dout + b * out_dim + pk * VN);
    if (scaled) {
#pragma unroll
      for (int j = 0; j < VN; ++j) g.v[j] = from_f32<T>(to_f32(g.v[j]) * scale);
    }
    for (int t = 0; t < bag; ++t) {
      long long const row = (long long)idx[b * bag + t];
      red_add_vec(dtable + row * out_dim + pk * VN, g);
    }
  }
}

template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    embed_bwd_scalar_kernel(TI const *__restrict__ idx, T const *__restrict__ dout,
                            T *__restrict__ dtable, int out_dim, int bag, long long batch,
                            double scale) {
  long long const total = batch * out_dim;
  long long const stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long const b = i / out_dim, off = i - b * out_dim;
    T g = (T)((double)dout[i] * scale);
    for (int t = 0; t < bag; ++t)
      atomic_add_any(dtable + (long long)idx[b * bag + t] * out_dim + off, g);
  }
}

template <typename TI>
__global__ void __launch_bounds__(kThreads) iota_mod_kernel(TI *p, size_t n, TI m) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = (TI)(i % (size_t)m);
}

// ------------------------------------------------------------------ gather
template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    gather_fwd_kernel(T const *__restrict__ in, TI const *__restrict__ index, T *__restrict__ out,
                      long long n, long long stride, long long in_dim, long long out_dim) {
  long long const gs = (long long)gridDim.x * blockDim.x;
  for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += gs) {
    long long outer = o / (stride * out_dim), rest = o % stride;
    out[o] = in[outer * stride * in_dim + (long long)index[o] * stride + rest];
  }
}
template <typename TI, typename T>
__global__ void __launch_bounds__(kThreads)
    gather_bwd_kernel(T const *__restrict__ dout, TI const *__restrict__ index,
                      T *__restrict__ din, long long n, long long stride, long long in_dim,
                      long long out_dim) {
  long long const gs = (long long)gridDim.x * blockDim.x;
  for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += gs) {
    long long outer = o / (stride * out_dim), rest = o % stride;
    atomic_add_any(din + outer * stride * in_dim + (long long)index[o] * stride + rest, dout[o]);
  }
}

// ------------------------------------------------------------------ top-k
// One CTA per row. The row sits in shared memory; k rounds of "block arg-max with
// lowest-index tie-break, then retire the winner". O(k * length / threads) per row.
__global__ void __launch_bounds__(kThreads)
    topk_fwd_kernel(float const *__restrict__ in, float *__restrict__ out,
                    int *__restrict__ indices, int length, int k) {
  extern __shared__ float row[];
  __shared__ float wv[kThreads / 32];
  __shared__ int wi[kThreads / 32];
  __shared__ int winner;
  float const *src = in + size_t(blockIdx.x) * length;
  for (int i = threadIdx.x; i < length; i += blockDim.x) row[i] = src[i];
  __syncthreads();
  int const lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int r = 0; r < k; ++r) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < length; i += blockDim.x) {
      float v = row[i];
      // retired slots hold NaN and never win; real NaNs are not ordered by the reference either
      if (v > bv || (v == bv && i < bi)) {
        bv = v;
        bi = i;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) {
        bv = ov;
        bi = oi;
      }
    }
    if (lane == 0) {
      wv[w] = bv;
      wi[w] = bi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      float fv = wv[0];
      int fi = wi[0];
      for (int j = 1; j < kThreads / 32; ++j)
        if (wv[j] > fv || (wv[j] == fv && wi[j] < fi)) {
          fv = wv[j];
          fi = wi[j];
        }
      if (fi == 0x7fffffff) {  // nothing left that compares (> -inf): take lowest live slot
        fi = 0;
        for (int i = 0; i < length; ++i)
          if (row[i] == row[i]) { fi = i; fv = row[i]; break; }
      }
      out[size_t(blockIdx.x) * k + r] = fv;
      indices[size_t(blockIdx.x) * k + r] = fi;
      winner = fi;
    }
    __syncthreads();
    if (threadIdx.x == 0) row[winner] = __int_as_float(0x7fc00000);  // retire
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads)
    topk_bwd_kernel(float const *__restrict__ dout, int const *__restrict__ indices,
                    float *__restrict__ din, size_t batch, int length, int k) {
  size_t const n = batch * size_t(k), gs = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += gs) {
    size_t b = i / k;
    // indices within a row are distinct, so plain read-modify-write is race free
    din[b * length + indices[i]] += dout[i];
  }
}

template <typename TI>
int embedding_fwd_t(cudaStream_t s, void const *input, void *output, void const *weight,
                    int dtype, int aggr, int in_dim, int out_dim, long long batch) {
  int bag = aggr == FFK_AGGR_NONE ? 1 : in_dim;
  bool aggregate = aggr != FFK_AGGR_NONE;
  double scale = aggr == FFK_AGGR_AVG ? 1.0 / double(bag) : 1.0;
  if (dtype == FFK_DOUBLE) {
    unsigned grid = stream_grid(size_t(batch) * out_dim, kThreads, 8);
    embed_fwd_scalar_kernel<TI, double><<<grid, kThreads, 0, s>>>(
        (TI const *)input, (double *)output, (double const *)weight, out_dim, bag, batch, scale,
        aggregate);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    if (out_dim % VN == 0 && aligned16(output) && aligned16(weight)) {
      int ppr = out_dim / VN;
      unsigned grid = stream_grid(size_t(batch) * ppr, kThreads, 8);
      embed_fwd_vec_kernel<TI, T><<<grid, kThreads, 0, s>>>((TI const *)input, (T *)output,
                                                            (T const *)weight, ppr, bag, batch,
                                                            (float)scale, aggregate);
    } else {
      unsigned grid = stream_grid(size_t(batch) * out_dim, kThreads, 8);
      if constexpr (std::is_same<T, float>::value) {
        embed_fwd_scalar_kernel<TI, float><<<grid, kThreads, 0, s>>>(
            (TI const *)input, (float *)output, (float const *)weight, out_dim, bag, batch, scale,
            aggregate);
      } else {
        set_error("embedding: 16-bit tables need out_dim %% 8 == 0 and 16-byte aligned tensors");
        return FFK_ERR_UNSUPPORTED;
      }
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

// `alpha` scales what is scattered: 1 for the gradient itself, -lr for the row-sparse SGD update
// (the table is the target then).
template <typename TI>
int embedding_bwd_t(cudaStream_t s, void const *input, void const *dout, void *dtable, int dtype,
                    int aggr, int in_dim, int out_dim, long long batch, double alpha = 1.0) {
  int bag = aggr == FFK_AGGR_NONE ? 1 : in_dim;
  double scale = (aggr == FFK_AGGR_AVG ? 1.0 / double(bag) : 1.0) * alpha;
  if (dtype == FFK_DOUBLE) {
    unsigned grid = stream_grid(size_t(batch) * out_dim, kThreads, 8);
    embed_bwd_scalar_kernel<TI, double><<<grid, kThreads, 0, s>>>(
        (TI const *)input, (double const *)dout, (double *)dtable, out_dim, bag, batch, scale);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    if (out_dim % VN == 0 && aligned16(dout) && aligned16(dtable)) {
      int ppr = out_dim / VN;
      unsigned grid = stream_grid(size_t(batch) * ppr, kThreads, 8);
      embed_bwd_vec_kernel<TI, T><<<grid, kThreads, 0, s>>>((TI const *)input, (T const *)dout,
                                                            (T *)dtable, ppr, bag, batch,
                                                            (float)scale, scale != 1.0);
    } else {
      if constexpr (std::is_same<T, float>::value) {
        unsigned grid = stream_grid(size_t(batch) * out_dim, kThreads, 8);
        embed_bwd_scalar_kernel<TI, float><<<grid, kThreads, 0, s>>>(
            (TI const *)input, (float const *)dout, (float *)dtable, out_dim, bag, batch, scale);
      } else {
        set_error("embedding: 16-bit tables need out_dim %% 8 == 0 and 16-byte aligned tensors");
        return FFK_ERR_UNSUPPORTED;
      }
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_embedding_forward(void *stream, void const *input, void *output, void const *weight,
                          int index_dtype, int dtype, int aggr, int in_dim, int out_dim,
                          int batch_size) {
  FFK_REQUIRE(out_dim > 0 && batch_size >= 0, "embedding: bad extents");
  FFK_REQUIRE(aggr >= FFK_AGGR_NONE && aggr <= FFK_AGGR_AVG, "embedding: bad aggr %d", aggr);
  FFK_REQUIRE(aggr == FFK_AGGR_NONE || in_dim > 0, "embedding: bag size must be positive");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output && weight, "embedding: null tensor");
  if (index_dtype == FFK_INT32)
    return embedding_fwd_t<int32_t>(as_stream(stream), input, output, weight, dtype, aggr, in_dim, out_dim, batch_size);
  if (index_dtype == FFK_INT64)
    return embedding_fwd_t<int64_t>(as_stream(stream), input, output, weight, dtype, aggr, in_dim, out_dim, batch_size);
  set_error("embedding: indices must be INT32 or INT64");
  return FFK_ERR_DTYPE;
}

int ffk_embedding_backward(void *stream, void const *input, void const *output_grad,
                           void *weight_grad, int index_dtype, int dtype, int aggr, int in_dim,
                           int out_dim, int batch_size) {
  FFK_REQUIRE(out_dim > 0 && batch_size >= 0, "embedding: bad extents");
  FFK_REQUIRE(aggr >= FFK_AGGR_NONE && aggr <= FFK_AGGR_AVG, "embedding: bad aggr %d", aggr);
  FFK_REQUIRE(aggr == FFK_AGGR_NONE || in_dim > 0, "embedding: bag size must be positive");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output_grad && weight_grad, "embedding backward: null tensor");
  if (index_dtype == FFK_INT32)
    return embedding_bwd_t<int32_t>(as_stream(stream), input, output_grad, weight_grad, dtype, aggr, in_dim, out_dim, batch_size);
  if (index_dtype == FFK_INT64)
    return embedding_bwd_t<int64_t>(as_stream(stream), input, output_grad, weight_grad, dtype, aggr, in_dim, out_dim, batch_size);
  set_error("embedding: indices must be INT32 or INT64");
  return FFK_ERR_DTYPE;
}

int ffk_embedding_sgd_update(void *stream, void const *input, void const *output_grad, void *weight,
                             float lr, int index_dtype, int dtype, int aggr, int in_dim, int out_dim,
                             int batch_size) {
  FFK_REQUIRE(out_dim > 0 && batch_size >= 0, "embedding: bad extents");
  FFK_REQUIRE(aggr >= FFK_AGGR_NONE && aggr <= FFK_AGGR_AVG, "embedding: bad aggr %d", aggr);
  FFK_REQUIRE(aggr == FFK_AGGR_NONE || in_dim > 0, "embedding: bag size must be positive");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output_grad && weight, "embedding update: null tensor");
  if (index_dtype == FFK_INT32)
    return embedding_bwd_t<int32_t>(as_stream(stream), input, output_grad, weight, dtype, aggr, in_dim, out_dim, batch_size, -double(lr));
  if (index_dtype == FFK_INT64)
    return embedding_bwd_t<int64_t>(as_stream(stream), input, output_grad, weight, dtype, aggr, in_dim, out_dim, batch_size, -double(lr));
  set_error("embedding: indices must be INT32 or INT64");
  return FFK_ERR_DTYPE;
}

int ffk_rand_generate_int(void *stream, void *ptr, size_t size, long long p, int index_dtype) {
  FFK_REQUIRE(p > 0, "rand_generate_int: modulus must be positive");
  if (size == 0) return FFK_OK;
  FFK_REQUIRE(ptr, "rand_generate_int: null pointer");
  unsigned grid = stream_grid(size, kThreads, 8);
  if (index_dtype == FFK_INT32)
    iota_mod_kernel<int32_t><<<grid, kThreads, 0, as_stream(stream)>>>((int32_t *)ptr, size, (int32_t)p);
  else if (index_dtype == FFK_INT64)
    iota_mod_kernel<int64_t><<<grid, kThreads, 0, as_stream(stream)>>>((int64_t *)ptr, size, (int64_t)p);
  else {
    set_error("rand_generate_int: dtype must be INT32 or INT64");
    return FFK_ERR_DTYPE;
  }
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

#define FFK_GATHER_DISPATCH(KERNEL, A, B, C)                                                     \
  do {                                                                                           \
    unsigned grid = stream_grid(size_t(output_size), kThreads, 8);                               \
    cudaStream_t s = as_stream(stream);                                                          \
    if (index_dtype != FFK_INT32 && index_dtype != FFK_INT64) {                                  \
      set_error("gather: indices must be INT32 or INT64");                                       \
      return FFK_ERR_DTYPE;                                                                      \
    }                                                                                            \
    bool i32 = index_dtype == FFK_INT32;                                                         \
    switch (dtype) {                                                                             \
      case FFK_FLOAT:                                                                            \
        if (i32) KERNEL<int32_t, float><<<grid, kThreads, 0, s>>>((float const *)A, (int32_t const *)B, (float *)C, output_size, stride, input_dim_size, output_dim_size); \
        else KERNEL<int64_t, float><<<grid, kThreads, 0, s>>>((float const *)A, (int64_t const *)B, (float *)C, output_size, stride, input_dim_size, output_dim_size); \
        break;                                                                                   \
      case FFK_BFLOAT16:                                                                         \
        if (i32) KERNEL<int32_t, __nv_bfloat16><<<grid, kThreads, 0, s>>>((__nv_bfloat16 const *)A, (int32_t const *)B, (__nv_bfloat16 *)C, output_size, stride, input_dim_size, output_dim_size); \
        else KERNEL<int64_t, __nv_bfloat16><<<grid, kThreads, 0, s>>>((__nv_bfloat16 const *)A, (int64_t const *)B, (__nv_bfloat16 *)C, output_size, stride, input_dim_size, output_dim_size); \
        break;                                                                                   \
      case FFK_HALF:                                                                             \
        if (i32) KERNEL<int32_t, __half><<<grid, kThreads, 0, s>>>((__half const *)A, (int32_t const *)B, (__half *)C, output_size, stride, input_dim_size, output_dim_size); \
        else KERNEL<int64_t, __half><<<grid, kThreads, 0, s>>>((__half const *)A, (int64_t const *)B, (__half *)C, output_size, stride, input_dim_size, output_dim_size); \
        break;                                                                                   \
      default:                                                                                   \
        set_error("gather: unsupported dtype %d", dtype);                                        \
        return FFK_ERR_DTYPE;                                                                    \
    }                                                                                            \
    FFK_LAUNCH_CHECK();                                                                          \
  } while (0)

int ffk_gather_forward(void *stream, void const *input, void const *index, void *output,
                       long long output_size, long long stride, long long input_dim_size,
                       long long output_dim_size, int index_dtype, int dtype) {
  FFK_REQUIRE(output_size >= 0 && stride > 0 && input_dim_size > 0 && output_dim_size > 0,
              "gather: bad extents");
  if (output_size == 0) return FFK_OK;
  FFK_REQUIRE(input && index && output, "gather: null tensor");
  FFK_GATHER_DISPATCH(gather_fwd_kernel, input, index, output);
  return FFK_OK;
}

int ffk_gather_backward(void *stream, void const *output_grad, void const *index,
                        void *input_grad, long long output_size, long long stride,
                        long long input_dim_size, long long output_dim_size, int index_dtype,
                        int dtype) {
  FFK_REQUIRE(output_size >= 0 && stride > 0 && input_dim_size > 0 && output_dim_size > 0,
              "gather: bad extents");
  if (output_size == 0) return FFK_OK;
  FFK_REQUIRE(output_grad && index && input_grad, "gather backward: null tensor");
  FFK_GATHER_DISPATCH(gather_bwd_kernel, output_grad, index, input_grad);
  return FFK_OK;
}

int ffk_topk_forward(void *stream, float const *input, float *output, int *indices,
                     size_t batch_size, int length, int k, int sorted) {
  (void)sorted;
  FFK_REQUIRE(length >= 0 && k >= 0 && k <= length, "topk: need 0 <= k <= length");
  if (batch_size == 0 || k == 0) return FFK_OK;
  FFK_REQUIRE(input && output && indices, "topk: null tensor");
  size_t smem = size_t(length) * sizeof(float);
  if (smem > size_t(device_info().max_smem_optin) - 1024) {
    set_error("topk: rows longer than %d floats are not supported", (device_info().max_smem_optin - 1024) / 4);
    return FFK_ERR_LIMIT;
  }
  FFK_CUDA(cudaFuncSetAttribute(topk_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  topk_fwd_kernel<<<(unsigned)batch_size, kThreads, smem, as_stream(stream)>>>(input, output, indices, length, k);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_topk_backward(void *stream, float const *out_grad, int const *indices, float *in_grad,
                      size_t batch_size, int length, int k) {
  FFK_REQUIRE(length >= 0 && k >= 0, "topk backward: bad extents");
  if (batch_size == 0 || k == 0) return FFK_OK;
  FFK_REQUIRE(out_grad && indices && in_grad, "topk backward: null tensor");
  unsigned grid = stream_grid(batch_size * size_t(k), kThreads, 8);
  topk_bwd_kernel<<<grid, kThreads, 0, as_stream(stream)>>>(out_grad, indices, in_grad, batch_size, length, k);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // extern "C"
