// HBM-bound pointwise kernels: ElementUnary, ElementBinary (with broadcasting), Cast,
// copy / accumulate / replica-sum (the bodies of the parallel ops' single-device forms),
// Dropout. All stream 16-byte packets with grid = SMs x resident CTAs.
#include "common.cuh"
#include "philox.cuh"

namespace ffk {
namespace {

constexpr int kThreads = 256;
constexpr int kResident = 8;  // CTAs per SM for streaming kernels (2048 threads / 256)

// ------------------------------------------------------------------ generic map kernels
// out[i] = f(a[i]) ; F is a functor on floats.
template <typename T, typename F>
__global__ void __launch_bounds__(kThreads) map1_kernel(T const *__restrict__ a, T *__restrict__ o,
                                                        size_t n, F f) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> x = ld_stream(a + i * VN), y;
#pragma unroll
    for (int j = 0; j < VN; ++j) y.v[j] = from_f32<T>(f(to_f32(x.v[j])));
    st_stream(o + i * VN, y);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride) o[i] = from_f32<T>(f(to_f32(a[i])));
}

// g[i] += f(x[i], y[i], dy[i])
template <typename T, typename F>
__global__ void __launch_bounds__(kThreads)
    unary_bwd_kernel(T const *__restrict__ x, T const *__restrict__ y, T const *__restrict__ dy,
                     T *__restrict__ g, size_t n, F f) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> xv = ld_stream(x + i * VN), yv = ld_stream(y + i * VN), dv = ld_stream(dy + i * VN);
    Vec16<T> gv = ld_vec(g + i * VN);
#pragma unroll
    for (int j = 0; j < VN; ++j)
      gv.v[j] = from_f32<T>(to_f32(gv.v[j]) + f(to_f32(xv.v[j]), to_f32(yv.v[j]), to_f32(dv.v[j])));
    st_vec(g + i * VN, gv);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride)
    g[i] = from_f32<T>(to_f32(g[i]) + f(to_f32(x[i]), to_f32(y[i]), to_f32(dy[i])));
}

struct UnaryFwd {
  int op;
  float s;
  __device__ __forceinline__ float operator()(float x) const {
    switch (op) {
      case FFK_OP_SCALAR_MULTIPLY: return x * s;
      case FFK_OP_SCALAR_ADD: return x + s;
      case FFK_OP_SCALAR_SUB: return x - s;
      case FFK_OP_SCALAR_TRUE_DIV: return x / s;
      case FFK_OP_POW: return powf(x, s);
      case FFK_OP_RELU: return fmaxf(x, 0.f);
      case FFK_OP_SIGMOID: return 1.f / (1.f + expf(-x));
      case FFK_OP_TANH: return tanhf(x);
      case FFK_OP_ELU: return x > 0.f ? x : s * (expf(x) - 1.f);
      case FFK_OP_EXP: return expf(x);
      case FFK_OP_GELU: return x * 0.5f * erfcf(-x * 0.70710678118654752f);
      case FFK_OP_RSQRT: return 1.0f / sqrtf(x);
      case FFK_OP_SIN: return sinf(x);
      case FFK_OP_COS: return cosf(x);
      default: return x;  // IDENTITY
    }
  }
};

struct UnaryBwd {
  int op;
  float s;
  // contribution to dx given x, y = f(x), dy
  __device__ __forceinline__ float operator()(float x, float y, float dy) const {
    switch (op) {
      case FFK_OP_SCALAR_MULTIPLY: return dy * s;
      case FFK_OP_SCALAR_ADD: return dy;
      case FFK_OP_SCALAR_SUB: return dy;
      case FFK_OP_SCALAR_TRUE_DIV: return dy / s;
      case FFK_OP_POW: return dy * s * powf(x, s - 1.f);
      case FFK_OP_RELU: return y > 0.f ? dy : 0.f;         // cudnnActivationBackward(RELU)
      case FFK_OP_SIGMOID: return dy * y * (1.f - y);
      case FFK_OP_TANH: return dy * (1.f - y * y);
      // cuDNN's rule, which the reference inherits: branch on the OUTPUT (y >= 0), so with the
      // reference's coefficient 0.0 the gradient is dy everywhere (tests/golden/cudnn_ops.npz)
      case FFK_OP_ELU: return y >= 0.f ? dy : dy * (y + s);
      case FFK_OP_EXP: return dy * expf(x);
      case FFK_OP_GELU:
        return dy * (0.5f * erfcf(-x * 0.70710678118654752f) +
                     x * 0.3989422804014327f * expf(-0.5f * x * x));
      case FFK_OP_RSQRT: return -0.5f * dy * y * y * y;
      case FFK_OP_SIN: return dy * cosf(x);
      case FFK_OP_COS: return -dy * sinf(x);
      default: return dy;  // IDENTITY
    }
  }
};

bool unary_supported(int op) {
  switch (op) {
    case FFK_OP_SCALAR_MULTIPLY: case FFK_OP_SCALAR_ADD: case FFK_OP_SCALAR_SUB:
    case FFK_OP_SCALAR_TRUE_DIV: case FFK_OP_POW: case FFK_OP_RELU: case FFK_OP_SIGMOID:
    case FFK_OP_TANH: case FFK_OP_ELU: case FFK_OP_EXP: case FFK_OP_GELU: case FFK_OP_RSQRT:
    case FFK_OP_SIN: case FFK_OP_COS: case FFK_OP_IDENTITY:
      return true;
    default: return false;
  }
}

// ------------------------------------------------------------------ binary
__device__ __forceinline__ float binary_fwd(int op, float l, float r) {
  switch (op) {
    case FFK_OP_EW_ADD: return l + r;
    case FFK_OP_EW_SUB: return l - r;
    case FFK_OP_EW_MUL: return l * r;
    case FFK_OP_EW_DIV: return l / r;
    case FFK_OP_EW_MAX: return fmaxf(l, r);
    default: return fminf(l, r);
  }
}
__device__ __forceinline__ float binary_dl(int op, float l, float r, float dy) {
  switch (op) {
    case FFK_OP_EW_ADD: case FFK_OP_EW_SUB: return dy;
    case FFK_OP_EW_MUL: return dy * r;
    case FFK_OP_EW_DIV: return dy / r;
    case FFK_OP_EW_MAX: return l >= r ? dy : 0.f;
    default: return l <= r ? dy : 0.f;
  }
}
__device__ __forceinline__ float binary_dr(int op, float l, float r, float dy) {
  switch (op) {
    case FFK_OP_EW_ADD: return dy;
    case FFK_OP_EW_SUB: return -dy;
    case FFK_OP_EW_MUL: return dy * l;
    case FFK_OP_EW_DIV: return -dy * l / (r * r);
    case FFK_OP_EW_MAX: return r >= l ? dy : 0.f;
    default: return r <= l ? dy : 0.f;
  }
}
bool binary_supported(int op) {
  return op == FFK_OP_EW_ADD || op == FFK_OP_EW_SUB || op == FFK_OP_EW_MUL ||
         op == FFK_OP_EW_DIV || op == FFK_OP_EW_MAX || op == FFK_OP_EW_MIN;
}
bool binary_needs_inputs(int op) { return !(op == FFK_OP_EW_ADD || op == FFK_OP_EW_SUB); }

template <typename T>
__global__ void __launch_bounds__(kThreads)
    binary_same_kernel(T const *__restrict__ l, T const *__restrict__ r, T *__restrict__ o,
                       size_t n, int op) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> a = ld_stream(l + i * VN), b = ld_stream(r + i * VN), y;
#pragma unroll
    for (int j = 0; j < VN; ++j) y.v[j] = from_f32<T>(binary_fwd(op, to_f32(a.v[j]), to_f32(b.v[j])));
    st_stream(o + i * VN, y);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride)
    o[i] = from_f32<T>(binary_fwd(op, to_f32(l[i]), to_f32(r[i])));
}

struct Bcast {
  int nd;
  long long out[5];
  long long ls[5];  // element strides of lhs per output dim (0 where broadcast)
  long long rs[5];
};

template <typename T>
__global__ void __launch_bounds__(kThreads)
    binary_bcast_kernel(T const *__restrict__ l, T const *__restrict__ r, T *__restrict__ o,
                        size_t n, int op, Bcast bc) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    size_t rem = i;
    long long lo = 0, ro = 0;
#pragma unroll
    for (int d = 4; d >= 0; --d) {
      if (d < bc.nd) {
        long long c = rem % bc.out[d];
        rem /= bc.out[d];
        lo += c * bc.ls[d];
        ro += c * bc.rs[d];
      }
    }
    o[i] = from_f32<T>(binary_fwd(op, to_f32(l[lo]), to_f32(r[ro])));
  }
}

// Same-shape backward: both grads accumulate in one pass over dY.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    binary_bwd_same_kernel(T const *__restrict__ dy, T const *__restrict__ l,
                           T const *__restrict__ r, T *__restrict__ dl, T *__restrict__ dr,
                           size_t n, int op, bool need_in) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> g = ld_stream(dy + i * VN), a, b;
    if (need_in) {
      a = ld_stream(l + i * VN);
      b = ld_stream(r + i * VN);
    }
    if (dl) {
      Vec16<T> t = ld_vec(dl + i * VN);
#pragma unroll
      for (int j = 0; j < VN; ++j)
        t.v[j] = from_f32<T>(to_f32(t.v[j]) + binary_dl(op, need_in ? to_f32(a.v[j]) : 0.f,
                                                        need_in ? to_f32(b.v[j]) : 0.f,
                                                        to_f32(g.v[j])));
      st_vec(dl + i * VN, t);
    }
    if (dr) {
      Vec16<T> t = ld_vec(dr + i * VN);
#pragma unroll
      for (int j = 0; j < VN; ++j)
        t.v[j] = from_f32<T>(to_f32(t.v[j]) + binary_dr(op, need_in ? to_f32(a.v[j]) : 0.f,
                                                        need_in ? to_f32(b.v[j]) : 0.f,
                                                        to_f32(g.v[j])));
      st_vec(dr + i * VN, t);
    }
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride) {
    float g = to_f32(dy[i]);
    float a = need_in ? to_f32(l[i]) : 0.f, b = need_in ? to_f32(r[i]) : 0.f;
    if (dl) dl[i] = from_f32<T>(to_f32(dl[i]) + binary_dl(op, a, b, g));
    if (dr) dr[i] = from_f32<T>(to_f32(dr[i]) + binary_dr(op, a, b, g));
  }
}

// Gradient of one side under broadcasting. One CTA per element of that side (grid-strided):
// the CTA sums the elementwise gradient over the output positions that map onto it.
// side: 0 = lhs, 1 = rhs. `tgt` strides select the side we are reducing for.
struct BcastBwd {
  int nd;
  long long out[5];
  long long ls[5], rs[5];
  long long tgt_dims[5];  // extents of the target side (1 where it broadcasts)
};

template <typename T>
__global__ void __launch_bounds__(kThreads)
    binary_bwd_bcast_kernel(T const *__restrict__ dy, T const *__restrict__ l,
                            T const *__restrict__ r, T *__restrict__ grad, size_t tgt_n,
                            size_t red_n, int op, int side, bool need_in, BcastBwd bc) {
  __shared__ float red[32];
  for (size_t e = blockIdx.x; e < tgt_n; e += gridDim.x) {
    // coordinates of the target element along non-broadcast dims
    long long base[5];
    size_t rem = e;
#pragma unroll
    for (int d = 4; d >= 0; --d) {
      if (d < bc.nd) {
        base[d] = rem % bc.tgt_dims[d];
        rem /= bc.tgt_dims[d];
      }
    }
    float acc = 0.f;
    for (size_t k = threadIdx.x; k < red_n; k += blockDim.x) {
      // k enumerates the broadcast dims (those where tgt extent is 1 and out > 1)
      size_t kr = k;
      long long oo = 0, lo = 0, ro = 0, ostride = 1;
#pragma unroll
      for (int d = 4; d >= 0; --d) {
        if (d < bc.nd) {
          long long c;
          if (bc.tgt_dims[d] == 1 && bc.out[d] > 1) {
            c = kr % bc.out[d];
            kr /= bc.out[d];
          } else {
            c = base[d];
          }
          oo += c * ostride;
          ostride *= bc.out[d];
          lo += c * bc.ls[d];
          ro += c * bc.rs[d];
        }
      }
      float g = to_f32(dy[oo]);
      float a = need_in ? to_f32(l[lo]) : 0.f, b = need_in ? to_f32(r[ro]) : 0.f;
      acc += side == 0 ? binary_dl(op, a, b, g) : binary_dr(op, a, b, g);
    }
    float tot = block_sum(acc, red);
    if (threadIdx.x == 0) grad[e] = from_f32<T>(to_f32(grad[e]) + tot);
    __syncthreads();
  }
}

// ------------------------------------------------------------------ cast
template <typename TI, typename TO>
__global__ void __launch_bounds__(kThreads)
    cast_kernel(TI const *__restrict__ in, TO *__restrict__ out, size_t n, bool accumulate) {
  size_t const stride = size_t(gridDim.x) * blockDim.x * 4;
  for (size_t i0 = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i0 < n; i0 += stride) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      size_t i = i0 + j;
      if (i < n) {
        TO v = static_cast<TO>(in[i]);
        out[i] = accumulate ? static_cast<TO>(v + out[i]) : v;
      }
    }
  }
}
// wrappers giving __half / bf16 the arithmetic the generic kernel needs
template <typename TI, typename TO> struct CastLaunch {
  static int run(void const *in, void *out, size_t n, bool acc, cudaStream_t s) {
    unsigned grid = stream_grid((n + 3) / 4, kThreads, kResident);
    cast_kernel<TI, TO><<<grid, kThreads, 0, s>>>(reinterpret_cast<TI const *>(in),
                                                  reinterpret_cast<TO *>(out), n, acc);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
};

// 16-bit floats go through float so that every pair has one conversion rule.
template <typename TI, typename TO>
__global__ void __launch_bounds__(kThreads)
    cast_via_float_kernel(TI const *__restrict__ in, TO *__restrict__ out, size_t n,
                          bool accumulate) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = to_f32(in[i]);
    if (accumulate) v += to_f32(out[i]);
    out[i] = from_f32<TO>(v);
  }
}

template <typename TO> int cast_from(int in_dt, void const *in, void *out, size_t n, bool acc,
                                     cudaStream_t s);

template <typename TI, typename TO>
int cast_dispatch2(void const *in, void *out, size_t n, bool acc, cudaStream_t s) {
  constexpr bool small_i = std::is_same<TI, __half>::value || std::is_same<TI, __nv_bfloat16>::value;
  constexpr bool small_o = std::is_same<TO, __half>::value || std::is_same<TO, __nv_bfloat16>::value;
  if constexpr (small_i || small_o) {
    if constexpr ((small_i || std::is_same<TI, float>::value) &&
                  (small_o || std::is_same<TO, float>::value)) {
      unsigned grid = stream_grid(n, kThreads, kResident);
      cast_via_float_kernel<TI, TO><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<TI const *>(in), reinterpret_cast<TO *>(out), n, acc);
      FFK_LAUNCH_CHECK();
      return FFK_OK;
    } else {
      set_error("cast: 16-bit floats convert to/from HALF, BFLOAT16 and FLOAT only");
      return FFK_ERR_DTYPE;
    }
  } else {
    return CastLaunch<TI, TO>::run(in, out, n, acc, s);
  }
}

template <typename TI>
int cast_dispatch1(int out_dt, void const *in, void *out, size_t n, bool acc, cudaStream_t s) {
  switch (out_dt) {
    case FFK_BOOL: return cast_dispatch2<TI, bool>(in, out, n, acc, s);
    case FFK_INT32: return cast_dispatch2<TI, int32_t>(in, out, n, acc, s);
    case FFK_INT64: return cast_dispatch2<TI, int64_t>(in, out, n, acc, s);
    case FFK_HALF: return cast_dispatch2<TI, __half>(in, out, n, acc, s);
    case FFK_FLOAT: return cast_dispatch2<TI, float>(in, out, n, acc, s);
    case FFK_DOUBLE: return cast_dispatch2<TI, double>(in, out, n, acc, s);
    case FFK_BFLOAT16: return cast_dispatch2<TI, __nv_bfloat16>(in, out, n, acc, s);
    default: set_error("cast: bad output dtype %d", out_dt); return FFK_ERR_DTYPE;
  }
}

int cast_any(void const *in, void *out, size_t n, int in_dt, int out_dt, bool acc,
             cudaStream_t s) {
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(in && out, "cast: null tensor");
  switch (in_dt) {
    case FFK_BOOL: return cast_dispatch1<bool>(out_dt, in, out, n, acc, s);
    case FFK_INT32: return cast_dispatch1<int32_t>(out_dt, in, out, n, acc, s);
    case FFK_INT64: return cast_dispatch1<int64_t>(out_dt, in, out, n, acc, s);
    case FFK_HALF: return cast_dispatch1<__half>(out_dt, in, out, n, acc, s);
    case FFK_FLOAT: return cast_dispatch1<float>(out_dt, in, out, n, acc, s);
    case FFK_DOUBLE: return cast_dispatch1<double>(out_dt, in, out, n, acc, s);
    case FFK_BFLOAT16: return cast_dispatch1<__nv_bfloat16>(out_dt, in, out, n, acc, s);
    default: set_error("cast: bad input dtype %d", in_dt); return FFK_ERR_DTYPE;
  }
}

// ------------------------------------------------------------------ copy / accumulate
// Bit-exact copy of `bytes` bytes in 16-byte packets (head/tail bytes singly).
__global__ void __launch_bounds__(kThreads)
    copy_bytes_kernel(uint8_t *__restrict__ dst, uint8_t const *__restrict__ src, size_t bytes) {
  size_t const nvec = bytes / 16, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  int4 const *s4 = reinterpret_cast<int4 const *>(src);
  int4 *d4 = reinterpret_cast<int4 *>(dst);
  // 4 packets in flight per thread
  size_t i = tid;
  for (; i + 3 * stride < nvec; i += 4 * stride) {
    int4 a = __ldcs(s4 + i), b = __ldcs(s4 + i + stride), c = __ldcs(s4 + i + 2 * stride),
         d = __ldcs(s4 + i + 3 * stride);
    __stcs(d4 + i, a);
    __stcs(d4 + i + stride, b);
    __stcs(d4 + i + 2 * stride, c);
    __stcs(d4 + i + 3 * stride, d);
  }
  for (; i < nvec; i += stride) __stcs(d4 + i, __ldcs(s4 + i));
  for (size_t b = nvec * 16 + tid; b < bytes; b += stride) dst[b] = src[b];
}
__global__ void __launch_bounds__(kThreads)
    copy_bytes_unaligned_kernel(uint8_t *__restrict__ dst, uint8_t const *__restrict__ src,
                                size_t bytes) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t b = size_t(blockIdx.x) * blockDim.x + threadIdx.x; b < bytes; b += stride)
    dst[b] = src[b];
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    accumulate_kernel(T *__restrict__ dst, T const *__restrict__ src, size_t n) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    Vec16<T> s = ld_stream(src + i * VN), d = ld_vec(dst + i * VN);
#pragma unroll
    for (int j = 0; j < VN; ++j) d.v[j] = from_f32<T>(to_f32(d.v[j]) + to_f32(s.v[j]));
    st_vec(dst + i * VN, d);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride)
    dst[i] = from_f32<T>(to_f32(dst[i]) + to_f32(src[i]));
}
template <typename T>
__global__ void __launch_bounds__(kThreads)
    accumulate_scalar_kernel(T *__restrict__ dst, T const *__restrict__ src, size_t n) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = dst[i] + src[i];
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    sum_replicas_kernel(T *__restrict__ dst, T const *__restrict__ src, size_t n, int reps,
                        bool accumulate, bool vec_ok) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = vec_ok ? n / VN : 0, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    float acc[VN];
    if (accumulate) {
      Vec16<T> d = ld_vec(dst + i * VN);
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] = to_f32(d.v[j]);
    } else {
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] = 0.f;
    }
    for (int rpl = 0; rpl < reps; ++rpl) {
      Vec16<T> s = ld_stream(src + size_t(rpl) * n + i * VN);
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] += to_f32(s.v[j]);
    }
    Vec16<T> o;
#pragma unroll
    for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>(acc[j]);
    st_vec(dst + i * VN, o);
  }
  for (size_t i = nvec * VN + tid; i < n; i += stride) {
    float acc = accumulate ? to_f32(dst[i]) : 0.f;
    for (int rpl = 0; rpl < reps; ++rpl) acc += to_f32(src[size_t(rpl) * n + i]);
    dst[i] = from_f32<T>(acc);
  }
}

// ------------------------------------------------------------------ dropout
// Philox4x32-10 and the per-element keep rule live in philox.cuh (shared with the GEMM epilogue).
// Fast path: one 16-byte packet per thread, coalesced; the VN keep-bits of a packet are
// OR-combined across the 32/VN lanes that share a mask word.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    dropout_fwd_kernel(T const *__restrict__ x, T *__restrict__ y, uint32_t *__restrict__ mask,
                       size_t nfull, uint32_t thresh16, bool drop_all, float scale, uint64_t seed,
                       uint64_t offset) {
  constexpr int VN = Vec16<T>::N;
  constexpr int LANES_PER_WORD = 32 / VN;
  size_t const nvec = nfull / VN, stride = size_t(gridDim.x) * blockDim.x;
  size_t const nvec_up = (nvec + 31) / 32 * 32;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec_up; i += stride) {
    bool const active = i < nvec;
    uint32_t bits = 0;
    if (active) {
      size_t const e0 = i * VN;
      if (!drop_all) {
        if constexpr (VN == 8) {
          bits = philox_keep8(e0 / 8, offset, seed, thresh16);
        } else {  // 4 elements: this packet's half of the 8-element block
          bits = (philox_keep8(e0 / 8, offset, seed, thresh16) >> (e0 % 8)) & 0xFu;
        }
      }
      Vec16<T> v = ld_stream(x + e0), o;
#pragma unroll
      for (int j = 0; j < VN; ++j)
        o.v[j] = from_f32<T>(((bits >> j) & 1u) ? to_f32(v.v[j]) * scale : 0.f);
      st_stream(y + e0, o);
    }
    uint32_t word = bits << ((i % LANES_PER_WORD) * VN);
#pragma unroll
    for (int o = 1; o < LANES_PER_WORD; o <<= 1) word |= __shfl_xor_sync(0xffffffffu, word, o);
    if (active && (i % LANES_PER_WORD) == 0) mask[i / LANES_PER_WORD] = word;
  }
}

// Generic path: one thread per mask word in [w_begin, w_end); any alignment, ragged tail.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    dropout_fwd_words_kernel(T const *__restrict__ x, T *__restrict__ y,
                             uint32_t *__restrict__ mask, size_t n, size_t w_begin, size_t w_end,
                             uint32_t thresh16, bool drop_all, float scale, uint64_t seed,
                             uint64_t offset) {
  size_t const stride = size_t(gridDim.x) * blockDim.x;
  for (size_t w = w_begin + size_t(blockIdx.x) * blockDim.x + threadIdx.x; w < w_end; w += stride) {
    uint32_t bits = 0;
    if (!drop_all) {
#pragma unroll
      for (int q = 0; q < 4; ++q) bits |= philox_keep8(w * 4 + q, offset, seed, thresh16) << (q * 8);
    }
    size_t const e0 = w * 32;
    if (n - e0 < 32) bits &= (1u << (n - e0)) - 1u;
    mask[w] = bits;
    for (int j = 0; j < 32 && e0 + j < n; ++j)
      y[e0 + j] = from_f32<T>(((bits >> j) & 1u) ? to_f32(x[e0 + j]) * scale : 0.f);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    dropout_bwd_kernel(T const *__restrict__ dy, T *__restrict__ dx,
                       uint32_t const *__restrict__ mask, size_t n, float scale, bool vec_ok) {
  constexpr int VN = Vec16<T>::N;
  size_t const nvec = vec_ok ? n / VN : 0, stride = size_t(gridDim.x) * blockDim.x;
  size_t const tid = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (size_t i = tid; i < nvec; i += stride) {
    size_t const e0 = i * VN;
    uint32_t const bits = __ldg(mask + e0 / 32) >> (e0 % 32);
    Vec16<T> v = ld_stream(dy + e0), o;
#pragma unroll
    for (int j = 0; j < VN; ++j)
      o.v[j] = from_f32<T>(((bits >> j) & 1u) ? to_f32(v.v[j]) * scale : 0.f);
    st_stream(dx + e0, o);
  }
  for (size_t e = nvec * VN + tid; e < n; e += stride)
    dx[e] = from_f32<T>(((mask[e / 32] >> (e % 32)) & 1u) ? to_f32(dy[e]) * scale : 0.f);
}

int fill_bcast(int ndims, long long const *ld, long long const *rd, long long const *od,
               long long (&out)[5], long long (&ls)[5], long long (&rs)[5], size_t &n_out,
               size_t &n_l, size_t &n_r) {
  FFK_REQUIRE(ndims >= 1 && ndims <= 5, "element_binary: ndims must be 1..5, got %d", ndims);
  FFK_REQUIRE(ld && rd && od, "element_binary: null shape array");
  n_out = n_l = n_r = 1;
  long long lstr = 1, rstr = 1;
  for (int d = ndims - 1; d >= 0; --d) {
    FFK_REQUIRE(od[d] > 0 && (ld[d] == od[d] || ld[d] == 1) && (rd[d] == od[d] || rd[d] == 1),
                "element_binary: dim %d not broadcastable (lhs %lld rhs %lld out %lld)", d, ld[d],
                rd[d], od[d]);
    out[d] = od[d];
    ls[d] = (ld[d] == 1 && od[d] != 1) ? 0 : lstr;
    rs[d] = (rd[d] == 1 && od[d] != 1) ? 0 : rstr;
    lstr *= ld[d];
    rstr *= rd[d];
    n_out *= od[d];
  }
  n_l = lstr;
  n_r = rstr;
  return FFK_OK;
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_element_unary_forward(void *stream, int op_type, float scalar, void const *input,
                              void *output, size_t n, int dtype) {
  FFK_REQUIRE(unary_supported(op_type), "element_unary: unsupported op_type %d", op_type);
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(input && output, "element_unary: null tensor");
  FFK_REQUIRE(aligned16(input) && aligned16(output), "element_unary: tensors must be 16-byte aligned");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
    map1_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), n,
        UnaryFwd{op_type, scalar});
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_element_unary_backward(void *stream, int op_type, float scalar, void const *input,
                               void *input_grad, void const *output, void const *output_grad,
                               size_t n, int dtype) {
  FFK_REQUIRE(unary_supported(op_type), "element_unary: unsupported op_type %d", op_type);
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(input && input_grad && output && output_grad, "element_unary backward: null tensor");
  FFK_REQUIRE(aligned16(input) && aligned16(input_grad) && aligned16(output) &&
                  aligned16(output_grad),
              "element_unary backward: tensors must be 16-byte aligned");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
    unary_bwd_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        reinterpret_cast<T const *>(input), reinterpret_cast<T const *>(output),
        reinterpret_cast<T const *>(output_grad), reinterpret_cast<T *>(input_grad), n,
        UnaryBwd{op_type, scalar});
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_element_binary_forward(void *stream, int op_type, void const *lhs, void const *rhs,
                               void *out, int ndims, long long const *lhs_dims_host,
                               long long const *rhs_dims_host, long long const *out_dims_host,
                               int dtype) {
  FFK_REQUIRE(binary_supported(op_type), "element_binary: unsupported op_type %d", op_type);
  cudaStream_t s = as_stream(stream);
  bool same = true;
  size_t n = 0, nl = 0, nr = 0;
  Bcast bc{};
  if (ndims == 0) {
    FFK_REQUIRE(out_dims_host, "element_binary: out_dims_host must hold the element count");
    n = size_t(out_dims_host[0]);
  } else {
    int rc = fill_bcast(ndims, lhs_dims_host, rhs_dims_host, out_dims_host, bc.out, bc.ls, bc.rs,
                        n, nl, nr);
    if (rc) return rc;
    bc.nd = ndims;
    same = (nl == n && nr == n);
  }
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(lhs && rhs && out, "element_binary: null tensor");
  if (same) {
    FFK_REQUIRE(aligned16(lhs) && aligned16(rhs) && aligned16(out),
                "element_binary: tensors must be 16-byte aligned");
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
      binary_same_kernel<T><<<grid, kThreads, 0, s>>>(reinterpret_cast<T const *>(lhs),
                                                      reinterpret_cast<T const *>(rhs),
                                                      reinterpret_cast<T *>(out), n, op_type);
    });
  } else {
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = stream_grid(n, kThreads, kResident);
      binary_bcast_kernel<T><<<grid, kThreads, 0, s>>>(reinterpret_cast<T const *>(lhs),
                                                       reinterpret_cast<T const *>(rhs),
                                                       reinterpret_cast<T *>(out), n, op_type, bc);
    });
  }
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_element_binary_backward(void *stream, int op_type, void const *out_grad, void const *lhs,
                                void const *rhs, void *lhs_grad, void *rhs_grad, int ndims,
                                long long const *lhs_dims_host, long long const *rhs_dims_host,
                                long long const *out_dims_host, int dtype) {
  FFK_REQUIRE(binary_supported(op_type), "element_binary: unsupported op_type %d", op_type);
  cudaStream_t s = as_stream(stream);
  size_t n = 0, nl = 0, nr = 0;
  BcastBwd bb{};
  bool same = true;
  if (ndims == 0) {
    FFK_REQUIRE(out_dims_host, "element_binary: out_dims_host must hold the element count");
    n = size_t(out_dims_host[0]);
  } else {
    int rc = fill_bcast(ndims, lhs_dims_host, rhs_dims_host, out_dims_host, bb.out, bb.ls, bb.rs,
                        n, nl, nr);
    if (rc) return rc;
    bb.nd = ndims;
    same = (nl == n && nr == n);
  }
  if (n == 0 || (!lhs_grad && !rhs_grad)) return FFK_OK;
  bool need_in = binary_needs_inputs(op_type);
  FFK_REQUIRE(out_grad, "element_binary backward: null out_grad");
  FFK_REQUIRE(!need_in || (lhs && rhs), "element_binary backward: op needs lhs and rhs");
  if (same) {
    FFK_REQUIRE(aligned16(out_grad) && aligned16(lhs_grad) && aligned16(rhs_grad) &&
                    (!need_in || (aligned16(lhs) && aligned16(rhs))),
                "element_binary backward: tensors must be 16-byte aligned");
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
      binary_bwd_same_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(out_grad), reinterpret_cast<T const *>(lhs),
          reinterpret_cast<T const *>(rhs), reinterpret_cast<T *>(lhs_grad),
          reinterpret_cast<T *>(rhs_grad), n, op_type, need_in);
    });
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  for (int side = 0; side < 2; ++side) {
    void *g = side == 0 ? lhs_grad : rhs_grad;
    if (!g) continue;
    long long const *td = side == 0 ? lhs_dims_host : rhs_dims_host;
    size_t tn = side == 0 ? nl : nr;
    for (int d = 0; d < ndims; ++d) bb.tgt_dims[d] = td[d];
    size_t red_n = n / tn;
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = (unsigned)std::min<size_t>(tn, size_t(device_info().sm_count) * kResident);
      binary_bwd_bcast_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(out_grad), reinterpret_cast<T const *>(lhs),
          reinterpret_cast<T const *>(rhs), reinterpret_cast<T *>(g), tn, red_n, op_type, side,
          need_in, bb);
    });
    FFK_LAUNCH_CHECK();
  }
  return FFK_OK;
}

int ffk_cast_forward(void *stream, void const *input, void *output, size_t n, int in_dtype,
                     int out_dtype) {
  return cast_any(input, output, n, in_dtype, out_dtype, false, as_stream(stream));
}
int ffk_cast_backward(void *stream, void const *src, void *dst, size_t n, int src_dtype,
                      int dst_dtype) {
  return cast_any(src, dst, n, src_dtype, dst_dtype, true, as_stream(stream));
}

int ffk_copy(void *stream, void *dst, void const *src, size_t n, int dtype) {
  size_t es = dtype_size(dtype);
  FFK_REQUIRE(es != 0, "copy: bad dtype %d", dtype);
  if (n == 0 || dst == src) return FFK_OK;
  FFK_REQUIRE(dst && src, "copy: null tensor");
  size_t bytes = n * es;
  cudaStream_t s = as_stream(stream);
  if (aligned16(dst) && aligned16(src)) {
    unsigned grid = stream_grid(bytes / 64 + 1, kThreads, kResident);
    copy_bytes_kernel<<<grid, kThreads, 0, s>>>(reinterpret_cast<uint8_t *>(dst),
                                                reinterpret_cast<uint8_t const *>(src), bytes);
  } else {
    unsigned grid = stream_grid(bytes, kThreads, kResident);
    copy_bytes_unaligned_kernel<<<grid, kThreads, 0, s>>>(
        reinterpret_cast<uint8_t *>(dst), reinterpret_cast<uint8_t const *>(src), bytes);
  }
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_accumulate(void *stream, void *dst, void const *src, size_t n, int dtype) {
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(dst && src, "accumulate: null tensor");
  cudaStream_t s = as_stream(stream);
  if (dtype == FFK_INT32 || dtype == FFK_INT64 || dtype == FFK_DOUBLE) {
    unsigned grid = stream_grid(n, kThreads, kResident);
    if (dtype == FFK_INT32)
      accumulate_scalar_kernel<int32_t><<<grid, kThreads, 0, s>>>((int32_t *)dst, (int32_t const *)src, n);
    else if (dtype == FFK_INT64)
      accumulate_scalar_kernel<int64_t><<<grid, kThreads, 0, s>>>((int64_t *)dst, (int64_t const *)src, n);
    else
      accumulate_scalar_kernel<double><<<grid, kThreads, 0, s>>>((double *)dst, (double const *)src, n);
    FFK_LAUNCH_CHECK();
    return FFK_OK;
  }
  FFK_REQUIRE(aligned16(dst) && aligned16(src), "accumulate: tensors must be 16-byte aligned");
  FFK_DISPATCH_FLOAT(dtype, T, {
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
    accumulate_kernel<T><<<grid, kThreads, 0, s>>>(reinterpret_cast<T *>(dst),
                                                   reinterpret_cast<T const *>(src), n);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_sum_replicas(void *stream, void *dst, void const *src, size_t n, int num_replicas,
                     int accumulate, int dtype) {
  FFK_REQUIRE(num_replicas >= 0, "sum_replicas: negative replica count");
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(dst && src, "sum_replicas: null tensor");
  FFK_DISPATCH_FLOAT(dtype, T, {
    bool vec_ok = aligned16(dst) && aligned16(src) && (n * sizeof(T)) % 16 == 0;
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
    sum_replicas_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        reinterpret_cast<T *>(dst), reinterpret_cast<T const *>(src), n, num_replicas,
        accumulate != 0, vec_ok);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

size_t ffk_dropout_mask_bytes(size_t n) { return ((n + 31) / 32) * sizeof(uint32_t); }

int ffk_dropout_forward(void *stream, void const *input, void *output, void *mask, size_t n,
                        float rate, unsigned long long seed, unsigned long long offset,
                        int dtype) {
  FFK_REQUIRE(rate >= 0.f && rate <= 1.f, "dropout: rate %f outside [0,1]", rate);
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(input && output && mask, "dropout: null tensor");
  float scale = rate < 1.f ? 1.f / (1.f - rate) : 0.f;
  bool drop_all = rate >= 1.f;
  uint32_t thresh = dropout_thresh16(rate);
  cudaStream_t s = as_stream(stream);
  size_t nwords = (n + 31) / 32;
  FFK_DISPATCH_FLOAT(dtype, T, {
    bool vec_ok = aligned16(input) && aligned16(output);
    size_t nfull = vec_ok ? n / 32 * 32 : 0;
    if (nfull) {
      unsigned grid = stream_grid(nfull / Vec16<T>::N, kThreads, kResident);
      dropout_fwd_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
          reinterpret_cast<uint32_t *>(mask), nfull, thresh, drop_all, scale, seed, offset);
    }
    if (nfull / 32 < nwords) {
      unsigned grid = stream_grid(nwords - nfull / 32, kThreads, kResident);
      dropout_fwd_words_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
          reinterpret_cast<uint32_t *>(mask), n, nfull / 32, nwords, thresh, drop_all, scale,
          seed, offset);
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_dropout_backward(void *stream, void const *output_grad, void *input_grad,
                         void const *mask, size_t n, float rate, int dtype) {
  FFK_REQUIRE(rate >= 0.f && rate <= 1.f, "dropout: rate %f outside [0,1]", rate);
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(output_grad && input_grad && mask, "dropout backward: null tensor");
  float scale = rate < 1.f ? 1.f / (1.f - rate) : 0.f;
  FFK_DISPATCH_FLOAT(dtype, T, {
    bool vec_ok = aligned16(output_grad) && aligned16(input_grad);
    unsigned grid = stream_grid(n / Vec16<T>::N + 1, kThreads, kResident);
    dropout_bwd_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
        reinterpret_cast<T const *>(output_grad), reinterpret_cast<T *>(input_grad),
        reinterpret_cast<uint32_t const *>(mask), n, scale, vec_ok);
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

}  // extern "C"
