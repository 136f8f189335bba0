// Softmax, LayerNorm and BatchNorm for sm_100a. Rows are read ONCE into registers as
// 16-byte packets (up to 8 packets per thread), reduced with warp shuffles and written
// back; rows that do not fit or are not 16-byte aligned take a generic looped kernel.
#include "common.cuh"
#include "ptx_sm100.cuh"

#include <cstdlib>
#include <mutex>

namespace ffk {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void block_sum2(float &a, float &b, float *red /* >= 64 floats */) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  a = warp_sum(a);
  b = warp_sum(b);
  __syncthreads();
  if (lane == 0) {
    red[w] = a;
    red[32 + w] = b;
  }
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  float ra = (lane < nw) ? red[lane] : 0.f, rb = (lane < nw) ? red[32 + lane] : 0.f;
  a = warp_sum(ra);
  b = warp_sum(rb);
}

inline int pick_p(long long nvec) {
  long long need = (nvec + kThreads - 1) / kThreads;
  if (need <= 1) return 1;
  if (need <= 2) return 2;
  if (need <= 4) return 4;
  if (need <= 8) return 8;
  return 0;
}

#define FFK_DISPATCH_P(p, P, ...)                                              \
  switch (p) {                                                                 \
    case 1: { constexpr int P = 1; __VA_ARGS__; break; }                       \
    case 2: { constexpr int P = 2; __VA_ARGS__; break; }                       \
    case 4: { constexpr int P = 4; __VA_ARGS__; break; }                       \
    default: { constexpr int P = 8; __VA_ARGS__; break; }                      \
  }

// ================================================================== softmax
template <typename T, int P>
__global__ void __launch_bounds__(kThreads)
    softmax_rows_kernel(T const *x, T *y, long long rows, int nvec) {
  constexpr int VN = Vec16<T>::N;
  __shared__ float red[32];
  bool const inplace = (x == y);  // attention normalises its score matrix in place
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *xr = x + r * (long long)nvec * VN;
    T *yr = y + r * (long long)nvec * VN;
    float v[P][VN];
    float mx = -INFINITY;
    // (64 resident warps per SM hide the latency here; issuing a thread's loads all at once
    //  measured slower for these long rows, unlike in the one-warp-per-row kernels below)
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * kThreads;
      if (i < nvec) {
        Vec16<T> t = inplace ? ld_vec(xr + i * VN) : ld_stream(xr + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          v[p][j] = to_f32(t.v[j]);
          mx = fmaxf(mx, v[p][j]);
        }
      }
    }
    mx = block_max(mx, red);
    float sum = 0.f;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * kThreads;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          v[p][j] = __expf(v[p][j] - mx);
          sum += v[p][j];
        }
      }
    }
    sum = block_sum(sum, red);
    float inv = 1.f / sum;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * kThreads;
      if (i < nvec) {
        Vec16<T> o;
#pragma unroll
        for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>(v[p][j] * inv);
        st_stream(yr + i * VN, o);
      }
    }
  }
}

// Short rows (<= 8 packets per lane): one WARP per row, no block barrier. A row of 512 bf16
// (the attention score rows) is 64 packets = 2 per lane.
template <typename T, int PW>
__global__ void __launch_bounds__(kThreads)
    softmax_warp_rows_kernel(T const *x, T *y, long long rows, int nvec) {
  constexpr int VN = Vec16<T>::N;
  int const lane = threadIdx.x & 31;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  bool const inplace = (x == y);
  for (long long r = warp0; r < rows; r += wstride) {
    T const *xr = x + r * (long long)nvec * VN;
    T *yr = y + r * (long long)nvec * VN;
    float v[PW][VN];
    float mx = -INFINITY;
    Vec16<T> t[PW];
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      int ii = i < nvec ? i : 0;
      t[p] = inplace ? ld_vec(xr + ii * VN) : ld_stream(xr + ii * VN);
    }
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          v[p][j] = to_f32(t[p].v[j]);
          mx = fmaxf(mx, v[p][j]);
        }
      }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          v[p][j] = __expf(v[p][j] - mx);
          sum += v[p][j];
        }
      }
    }
    sum = warp_sum(sum);
    float inv = 1.f / sum;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        Vec16<T> o;
#pragma unroll
        for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>(v[p][j] * inv);
        st_stream(yr + i * VN, o);
      }
    }
  }
}

// dx = y (.) (dy - sum(dy (.) y)) for contiguous short rows, one warp per row; dx may alias dy.
template <typename T, int PW>
__global__ void __launch_bounds__(kThreads)
    softmax_jacobian_warp_rows_kernel(T *dx, T const *dy, T const *__restrict__ y, long long rows,
                                      int nvec) {
  constexpr int VN = Vec16<T>::N;
  int const lane = threadIdx.x & 31;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  for (long long r = warp0; r < rows; r += wstride) {
    long long const base = r * (long long)nvec * VN;
    float g[PW][VN], p_[PW][VN];
    float dot = 0.f;
    Vec16<T> a[PW], b[PW];
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      int ii = i < nvec ? i : 0;
      a[p] = ld_vec(dy + base + ii * VN);
      b[p] = ld_stream(y + base + ii * VN);
    }
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          g[p][j] = to_f32(a[p].v[j]);
          p_[p][j] = to_f32(b[p].v[j]);
          dot += g[p][j] * p_[p][j];
        }
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        Vec16<T> o;
#pragma unroll
        for (int j = 0; j < VN; ++j) o.v[j] = from_f32<T>(p_[p][j] * (g[p][j] - dot));
        st_vec(dx + base + i * VN, o);
      }
    }
  }
}

inline int pick_pw(long long nvec) {
  long long need = (nvec + 31) / 32;
  if (need <= 1) return 1;
  if (need <= 2) return 2;
  if (need <= 4) return 4;
  if (need <= 8) return 8;
  return 0;
}

// Contiguous rows that cannot be read in 16-byte packets (30522 columns of bf16: the rows are only
// 4-byte aligned) but fit shared memory: one CTA per row, the row is brought in once element by
// element (coalesced), reduced twice from shared memory, and written once.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    softmax_rows_smem_kernel(T const *x, T *y, long long rows, long long dim) {
  extern __shared__ __align__(16) uint8_t sm_row_raw[];
  T *row = reinterpret_cast<T *>(sm_row_raw);
  __shared__ float red[32];
  // 16-bit rows of even length whose base is 4-byte aligned move two elements per access
  bool const pairs = sizeof(T) == 2 && dim % 2 == 0 &&
                     (reinterpret_cast<uintptr_t>(x) & 3) == 0 && (reinterpret_cast<uintptr_t>(y) & 3) == 0;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *xr = x + r * dim;
    T *yr = y + r * dim;
    float mx = -INFINITY;
    if (pairs) {
      struct alignas(4) Pair { T a, b; };
      Pair const *xp = reinterpret_cast<Pair const *>(xr);
      Pair *yp = reinterpret_cast<Pair *>(yr);
      Pair *rp = reinterpret_cast<Pair *>(row);
      long long const np = dim / 2;
      for (long long i = threadIdx.x; i < np; i += kThreads) {
        Pair const t = xp[i];
        rp[i] = t;
        mx = fmaxf(mx, fmaxf(to_f32(t.a), to_f32(t.b)));
      }
      mx = block_max(mx, red);
      float sum = 0.f;
      for (long long i = threadIdx.x; i < np; i += kThreads) {
        Pair const t = rp[i];
        sum += __expf(to_f32(t.a) - mx) + __expf(to_f32(t.b) - mx);
      }
      sum = block_sum(sum, red);
      float const inv = 1.f / sum;
      for (long long i = threadIdx.x; i < np; i += kThreads) {
        Pair const t = rp[i];
        Pair o;
        o.a = from_f32<T>(__expf(to_f32(t.a) - mx) * inv);
        o.b = from_f32<T>(__expf(to_f32(t.b) - mx) * inv);
        yp[i] = o;
      }
      __syncthreads();
      continue;
    }
    for (long long i = threadIdx.x; i < dim; i += kThreads) {
      T const t = xr[i];
      row[i] = t;
      mx = fmaxf(mx, to_f32(t));
    }
    mx = block_max(mx, red);
    float sum = 0.f;
    for (long long i = threadIdx.x; i < dim; i += kThreads) sum += __expf(to_f32(row[i]) - mx);
    sum = block_sum(sum, red);
    float const inv = 1.f / sum;
    for (long long i = threadIdx.x; i < dim; i += kThreads)
      yr[i] = from_f32<T>(__expf(to_f32(row[i]) - mx) * inv);
    __syncthreads();  // the row buffer is reused by the next row
  }
}

// Generic: one warp per (outer, inner) line, element stride `inner`.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    softmax_lines_kernel(T const *x, T *y, long long outer,
                         long long dim, long long inner) {
  long long const lines = outer * inner;
  int const lane = threadIdx.x & 31;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  for (long long l = warp0; l < lines; l += wstride) {
    long long o = l / inner, in = l - o * inner;
    T const *xl = x + o * dim * inner + in;
    T *yl = y + o * dim * inner + in;
    float mx = -INFINITY;
    for (long long k = lane; k < dim; k += 32) mx = fmaxf(mx, to_f32(xl[k * inner]));
    mx = warp_max(mx);
    float sum = 0.f;
    for (long long k = lane; k < dim; k += 32) sum += __expf(to_f32(xl[k * inner]) - mx);
    sum = warp_sum(sum);
    float inv = 1.f / sum;
    for (long long k = lane; k < dim; k += 32)
      yl[k * inner] = from_f32<T>(__expf(to_f32(xl[k * inner]) - mx) * inv);
  }
}

// Strided case with a wide inner extent: one thread per (outer, inner) line so that
// consecutive threads touch consecutive addresses.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    softmax_strided_kernel(T const *x, T *y, long long outer,
                           long long dim, long long inner) {
  long long const lines = outer * inner;
  long long const stride = (long long)gridDim.x * blockDim.x;
  for (long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x; l < lines; l += stride) {
    long long o = l / inner, in = l - o * inner;
    T const *xl = x + o * dim * inner + in;
    T *yl = y + o * dim * inner + in;
    float mx = -INFINITY;
    for (long long k = 0; k < dim; ++k) mx = fmaxf(mx, to_f32(xl[k * inner]));
    float sum = 0.f;
    for (long long k = 0; k < dim; ++k) sum += __expf(to_f32(xl[k * inner]) - mx);
    float inv = 1.f / sum;
    for (long long k = 0; k < dim; ++k)
      yl[k * inner] = from_f32<T>(__expf(to_f32(xl[k * inner]) - mx) * inv);
  }
}

// dx = y * (dy - sum(dy*y)) along dim; one warp per line (generic strides).
template <typename T>
__global__ void __launch_bounds__(kThreads)
    softmax_jacobian_kernel(T *dx, T const *dy, T const *__restrict__ y,
                            long long outer, long long dim, long long inner) {
  long long const lines = outer * inner;
  int const lane = threadIdx.x & 31;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  for (long long l = warp0; l < lines; l += wstride) {
    long long o = l / inner, in = l - o * inner;
    long long base = o * dim * inner + in;
    float dot = 0.f;
    for (long long k = lane; k < dim; k += 32)
      dot += to_f32(dy[base + k * inner]) * to_f32(y[base + k * inner]);
    dot = warp_sum(dot);
    for (long long k = lane; k < dim; k += 32) {
      long long i = base + k * inner;
      dx[i] = from_f32<T>(to_f32(y[i]) * (to_f32(dy[i]) - dot));
    }
  }
}

// ================================================================== layer norm
template <typename T, int P>
__global__ void __launch_bounds__(kThreads)
    layer_norm_fwd_rows_kernel(T const *__restrict__ x, T *__restrict__ y,
                               T const *__restrict__ gamma, T const *__restrict__ beta,
                               float *__restrict__ mean, float *__restrict__ rstd, long long rows,
                               int nvec, float eps) {
  constexpr int VN = Vec16<T>::N;
  __shared__ float red[64];
  float const invn = 1.f / float(nvec * VN);
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *xr = x + r * (long long)nvec * VN;
    T *yr = y + r * (long long)nvec * VN;
    float v[P][VN];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * kThreads;
      if (i < nvec) {
        Vec16<T> t = ld_stream(xr + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          v[p][j] = to_f32(t.v[j]);
          s1 += v[p][j];
          s2 += v[p][j] * v[p][j];
        }
      }
    }
    block_sum2(s1, s2, red);
    float mu = s1 * invn;
    float var = fmaxf(s2 * invn - mu * mu, 0.f);
    float rs = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      mean[r] = mu;
      rstd[r] = rs;
    }
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * kThreads;
      if (i < nvec) {
        Vec16<T> o, g, b;
        if (gamma) g = ld_vec(gamma + i * VN);
        if (beta) b = ld_vec(beta + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float u = (v[p][j] - mu) * rs;
          if (gamma) u *= to_f32(g.v[j]);
          if (beta) u += to_f32(b.v[j]);
          o.v[j] = from_f32<T>(u);
        }
        st_stream(yr + i * VN, o);
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    layer_norm_fwd_generic_kernel(T const *__restrict__ x, T *__restrict__ y,
                                  T const *__restrict__ gamma, T const *__restrict__ beta,
                                  float *__restrict__ mean, float *__restrict__ rstd,
                                  long long rows, long long N, float eps) {
  __shared__ float red[64];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *xr = x + r * N;
    T *yr = y + r * N;
    float s1 = 0.f, s2 = 0.f;
    for (long long j = threadIdx.x; j < N; j += blockDim.x) {
      float t = to_f32(xr[j]);
      s1 += t;
      s2 += t * t;
    }
    block_sum2(s1, s2, red);
    float mu = s1 / float(N);
    float var = fmaxf(s2 / float(N) - mu * mu, 0.f);
    float rs = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      mean[r] = mu;
      rstd[r] = rs;
    }
    for (long long j = threadIdx.x; j < N; j += blockDim.x) {
      float t = (to_f32(xr[j]) - mu) * rs;
      if (gamma) t *= to_f32(gamma[j]);
      if (beta) t += to_f32(beta[j]);
      yr[j] = from_f32<T>(t);
    }
  }
}

// Fused backward: one pass over dY and X gives dX (accumulated) and, when `part` is
// non-null, this CTA's partial column sums for dgamma / dbeta:
//   part[(blockIdx.x*2 + 0)*N + j] = sum over this CTA's rows of dy*(x-mean)*rstd
//   part[(blockIdx.x*2 + 1)*N + j] = sum over this CTA's rows of dy
// gamma and the two running column sums live in shared memory (3*N floats, accessed as
// float4 so a warp touches consecutive banks); each thread owns fixed columns, so no
// synchronisation is needed on them. Registers hold only the current row's dY and X
// packets, which keeps 3-4 CTAs resident per SM to overlap one row's reduction with the
// next row's loads.
template <typename T, int P, int TPB>
__global__ void __launch_bounds__(TPB)
    layer_norm_bwd_rows_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                               T *__restrict__ dx, T const *__restrict__ gamma,
                               float const *__restrict__ mean, float const *__restrict__ rstd,
                               float *__restrict__ part, long long rows, int nvec, bool dx_acc) {
  constexpr int VN = Vec16<T>::N;
  constexpr int Q = VN / 4;  // float4 groups per packet
  extern __shared__ float4 ln_smem[];
  __shared__ float red[64];
  long long const N = (long long)nvec * VN;
  float4 *sg = ln_smem;                 // gamma
  float4 *spg = ln_smem + N / 4;        // running sum for dgamma
  float4 *spb = ln_smem + 2 * (N / 4);  // running sum for dbeta
  float const invn = 1.f / float(N);
#pragma unroll
  for (int p = 0; p < P; ++p) {
    int i = threadIdx.x + p * TPB;
    if (i < nvec) {
      float gv[VN];
#pragma unroll
      for (int j = 0; j < VN; ++j) gv[j] = 1.f;
      if (gamma) {
        Vec16<T> t = ld_vec(gamma + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) gv[j] = to_f32(t.v[j]);
      }
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        sg[i * Q + q] = make_float4(gv[q * 4], gv[q * 4 + 1], gv[q * 4 + 2], gv[q * 4 + 3]);
        spg[i * Q + q] = make_float4(0.f, 0.f, 0.f, 0.f);
        spb[i * Q + q] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  // no barrier needed: every thread only ever touches the shared slots it just wrote
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *dyr = dy + r * N;
    T const *xr = x + r * N;
    float dv[P][VN], xv[P][VN];
    float ds = 0.f, db = 0.f;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * TPB;
      if (i < nvec) {
        Vec16<T> a = ld_stream(dyr + i * VN), b = ld_stream(xr + i * VN);
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          float4 g4 = sg[i * Q + q];
          float gq[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float d = to_f32(a.v[q * 4 + j]), xx = to_f32(b.v[q * 4 + j]);
            dv[p][q * 4 + j] = d;
            xv[p][q * 4 + j] = xx;
            ds += d * xx * gq[j];
            db += d * gq[j];
          }
        }
      }
    }
    block_sum2(ds, db, red);
    float const mu = mean[r], rs = rstd[r];
    float const c1 = (db * mu - ds) * rs * rs * rs * invn;
    float const c2 = -(c1 * mu + db * rs * invn);
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * TPB;
      if (i < nvec) {
        if (dx) {
          T *dxr = dx + r * N;
          Vec16<T> o = Vec16<T>{};  // dx_acc: dX += ..., else dX = ... (no read of the old value)
          if (dx_acc) o = ld_vec(dxr + i * VN);
#pragma unroll
          for (int q = 0; q < Q; ++q) {
            float4 g4 = sg[i * Q + q];
            float gq[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
              o.v[q * 4 + j] = from_f32<T>(to_f32(o.v[q * 4 + j]) + rs * dv[p][q * 4 + j] * gq[j] +
                                           c1 * xv[p][q * 4 + j] + c2);
          }
          st_vec(dxr + i * VN, o);
        }
        if (part) {
#pragma unroll
          for (int q = 0; q < Q; ++q) {
            float4 a4 = spg[i * Q + q], b4 = spb[i * Q + q];
            a4.x += dv[p][q * 4 + 0] * (xv[p][q * 4 + 0] - mu) * rs;
            a4.y += dv[p][q * 4 + 1] * (xv[p][q * 4 + 1] - mu) * rs;
            a4.z += dv[p][q * 4 + 2] * (xv[p][q * 4 + 2] - mu) * rs;
            a4.w += dv[p][q * 4 + 3] * (xv[p][q * 4 + 3] - mu) * rs;
            b4.x += dv[p][q * 4 + 0];
            b4.y += dv[p][q * 4 + 1];
            b4.z += dv[p][q * 4 + 2];
            b4.w += dv[p][q * 4 + 3];
            spg[i * Q + q] = a4;
            spb[i * Q + q] = b4;
          }
        }
      }
    }
  }
  if (part) {
    float4 *pgr = reinterpret_cast<float4 *>(part + (size_t(blockIdx.x) * 2 + 0) * N);
    float4 *pbr = reinterpret_cast<float4 *>(part + (size_t(blockIdx.x) * 2 + 1) * N);
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * TPB;
      if (i < nvec) {
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          pgr[i * Q + q] = spg[i * Q + q];
          pbr[i * Q + q] = spb[i * Q + q];
        }
      }
    }
  }
}

// Sum the per-CTA partials (deterministic order) and write / accumulate. A CTA owns 32
// columns; its 32 row-lanes stride over the partial slabs four loads deep, then combine in
// lane order through shared memory.
template <typename T>
__global__ void __launch_bounds__(1024)
    layer_norm_param_reduce_kernel(float const *__restrict__ part, int nparts, long long N,
                                   T *__restrict__ dgamma, T *__restrict__ dbeta,
                                   bool accumulate) {
  __shared__ float sg[32][33], sb[32][33];
  int const cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  long long const j = (long long)blockIdx.x * 32 + cx;
  float ag = 0.f, ab = 0.f;
  if (j < N) {
    int p = ry;
    for (; p + 96 < nparts; p += 128) {
      float g0 = part[(size_t(p) * 2 + 0) * N + j], b0 = part[(size_t(p) * 2 + 1) * N + j];
      float g1 = part[(size_t(p + 32) * 2 + 0) * N + j], b1 = part[(size_t(p + 32) * 2 + 1) * N + j];
      float g2 = part[(size_t(p + 64) * 2 + 0) * N + j], b2 = part[(size_t(p + 64) * 2 + 1) * N + j];
      float g3 = part[(size_t(p + 96) * 2 + 0) * N + j], b3 = part[(size_t(p + 96) * 2 + 1) * N + j];
      ag += (g0 + g1) + (g2 + g3);
      ab += (b0 + b1) + (b2 + b3);
    }
    for (; p < nparts; p += 32) {
      ag += part[(size_t(p) * 2 + 0) * N + j];
      ab += part[(size_t(p) * 2 + 1) * N + j];
    }
  }
  sg[ry][cx] = ag;
  sb[ry][cx] = ab;
  __syncthreads();
  if (ry == 0 && j < N) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      tg += sg[i][cx];
      tb += sb[i][cx];
    }
    if (dgamma) dgamma[j] = from_f32<T>(accumulate ? to_f32(dgamma[j]) + tg : tg);
    if (dbeta) dbeta[j] = from_f32<T>(accumulate ? to_f32(dbeta[j]) + tb : tb);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
    layer_norm_bwd_generic_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                                  T *__restrict__ dx, T const *__restrict__ gamma,
                                  float const *__restrict__ mean, float const *__restrict__ rstd,
                                  long long rows, long long N, bool dx_acc) {
  __shared__ float red[64];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    T const *dyr = dy + r * N;
    T const *xr = x + r * N;
    float ds = 0.f, db = 0.f;
    for (long long j = threadIdx.x; j < N; j += blockDim.x) {
      float gv = gamma ? to_f32(gamma[j]) : 1.f;
      float d = to_f32(dyr[j]);
      ds += d * to_f32(xr[j]) * gv;
      db += d * gv;
    }
    block_sum2(ds, db, red);
    float const mu = mean[r], rs = rstd[r], invn = 1.f / float(N);
    float const c1 = (db * mu - ds) * rs * rs * rs * invn;
    float const c2 = -(c1 * mu + db * rs * invn);
    T *dxr = dx + r * N;
    for (long long j = threadIdx.x; j < N; j += blockDim.x) {
      float gv = gamma ? to_f32(gamma[j]) : 1.f;
      dxr[j] = from_f32<T>((dx_acc ? to_f32(dxr[j]) : 0.f) + rs * to_f32(dyr[j]) * gv + c1 * to_f32(xr[j]) + c2);
    }
  }
}

// Column reduction for dgamma / dbeta without scratch: a CTA owns 32 columns, 8 row-lanes.
template <typename T>
__global__ void __launch_bounds__(kThreads)
    layer_norm_param_columns_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                                    float const *__restrict__ mean,
                                    float const *__restrict__ rstd, T *__restrict__ dgamma,
                                    T *__restrict__ dbeta, long long rows, long long N,
                                    bool accumulate) {
  __shared__ float sg[8][33], sb[8][33];
  int const cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  long long const col = (long long)blockIdx.x * 32 + cx;
  float ag = 0.f, ab = 0.f;
  if (col < N) {
    for (long long r = ry; r < rows; r += 8) {
      float d = to_f32(dy[r * N + col]);
      ag += d * (to_f32(x[r * N + col]) - mean[r]) * rstd[r];
      ab += d;
    }
  }
  sg[ry][cx] = ag;
  sb[ry][cx] = ab;
  __syncthreads();
  if (ry == 0 && col < N) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      tg += sg[i][cx];
      tb += sb[i][cx];
    }
    if (dgamma) dgamma[col] = from_f32<T>(accumulate ? to_f32(dgamma[col]) + tg : tg);
    if (dbeta) dbeta[col] = from_f32<T>(accumulate ? to_f32(dbeta[col]) + tb : tb);
  }
}

// Short rows: one warp per row for LayerNorm as well (a d_model = 1024 bf16 row is 128 packets).
template <typename T, int PW>
__global__ void __launch_bounds__(kThreads, (PW <= 2 ? 3 : PW <= 4 ? 2 : 1))
    layer_norm_fwd_warp_rows_kernel(T const *__restrict__ x, T *__restrict__ y,
                                    T const *__restrict__ gamma, T const *__restrict__ beta,
                                    float *__restrict__ mean, float *__restrict__ rstd,
                                    long long rows, int nvec, float eps) {
  constexpr int VN = Vec16<T>::N;
  int const lane = threadIdx.x & 31;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  float const invn = 1.f / float(nvec * VN);
  // gamma / beta of this lane's columns stay packed in registers for every row the warp visits
  Vec16<T> gk[PW], bk[PW];
#pragma unroll
  for (int p = 0; p < PW; ++p) {
    int i = lane + p * 32;
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      gk[p].v[j] = from_f32<T>(1.f);
      bk[p].v[j] = from_f32<T>(0.f);
    }
    if (gamma && i < nvec) gk[p] = ld_vec(gamma + i * VN);
    if (beta && i < nvec) bk[p] = ld_vec(beta + i * VN);
  }
  // the row stays PACKED in registers (widened where used): 4 registers per packet instead of VN.
  // The NEXT row's packets are requested before this row is reduced, so that a warp always has
  // two rows of loads in flight.
  Vec16<T> t[PW], tn[PW];
  if (warp0 < rows) {
    T const *xr = x + warp0 * (long long)nvec * VN;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      t[p] = ld_stream(xr + (i < nvec ? i : 0) * VN);
    }
  }
  for (long long r = warp0; r < rows; r += wstride) {
    T *yr = y + r * (long long)nvec * VN;
    long long const rn = r + wstride;
    if (rn < rows) {
      T const *xn = x + rn * (long long)nvec * VN;
#pragma unroll
      for (int p = 0; p < PW; ++p) {
        int i = lane + p * 32;
        tn[p] = ld_stream(xn + (i < nvec ? i : 0) * VN);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      if (lane + p * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float const v = to_f32(t[p].v[j]);
          s1 += v;
          s2 += v * v;
        }
      }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    float mu = s1 * invn;
    float var = fmaxf(s2 * invn - mu * mu, 0.f);
    float rs = rsqrtf(var + eps);
    if (lane == 0) {
      mean[r] = mu;
      rstd[r] = rs;
    }
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        Vec16<T> o;
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float u = (to_f32(t[p].v[j]) - mu) * rs;
          u = u * to_f32(gk[p].v[j]) + to_f32(bk[p].v[j]);
          o.v[j] = from_f32<T>(u);
        }
        st_stream(yr + i * VN, o);
      }
    }
#pragma unroll
    for (int p = 0; p < PW; ++p) t[p] = tn[p];
  }
}

// Forward for long rows through a bulk-copy ring: three rows in flight per CTA, the row is read
// from shared memory twice (statistics, then normalisation), so nothing of it lives in registers
// and four CTAs stay resident per SM at 16 KB rows.
template <typename T, int P>
__global__ void __launch_bounds__(256)
    layer_norm_fwd_cta_ring_kernel(T const *__restrict__ x, T *__restrict__ y,
                                   T const *__restrict__ gamma, T const *__restrict__ beta,
                                   float *__restrict__ mean, float *__restrict__ rstd,
                                   long long rows, int nvec, float eps) {
  constexpr int VN = Vec16<T>::N, STAGES = 3;
  extern __shared__ __align__(128) uint8_t ring_smem[];
  __shared__ float red[64];
  __shared__ __align__(8) uint64_t bars[STAGES];
  long long const N = (long long)nvec * VN;
  uint32_t const row_bytes = uint32_t(N * sizeof(T));
  if (threadIdx.x == 0) {
    for (int s2 = 0; s2 < STAGES; ++s2) ptx::mbar_init(&bars[s2], 1);
    ptx::fence_mbar_init();
    for (int s2 = 0; s2 < STAGES; ++s2) {
      long long r = blockIdx.x + (long long)s2 * gridDim.x;
      if (r < rows) {
        ptx::mbar_arrive_expect_tx(&bars[s2], row_bytes);
        ptx::bulk_load(ring_smem + size_t(s2) * row_bytes, x + r * N, row_bytes, &bars[s2]);
      }
    }
  }
  __syncthreads();
  float const invn = 1.f / float(N);
  // gamma / beta of this thread's (at most P) packets, packed, for every row the CTA visits: the
  // normalisation pass then touches shared memory only (they used to be re-read through L1 per row,
  // which is where the pass stalled)
  Vec16<T> gk[P], bk[P];
#pragma unroll
  for (int p = 0; p < P; ++p) {
    int const i = threadIdx.x + p * 256;
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      gk[p].v[j] = from_f32<T>(1.f);
      bk[p].v[j] = from_f32<T>(0.f);
    }
    if (gamma && i < nvec) gk[p] = ld_vec(gamma + i * VN);
    if (beta && i < nvec) bk[p] = ld_vec(beta + i * VN);
  }
  int st = 0;
  uint32_t phase = 0;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    ptx::mbar_wait(&bars[st], phase);
    uint8_t const *src = ring_smem + size_t(st) * row_bytes;
    float s1 = 0.f, s2 = 0.f;
    for (int i = threadIdx.x; i < nvec; i += 256) {
      Vec16<T> t = *reinterpret_cast<Vec16<T> const *>(src + size_t(i) * 16);
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        float const v = to_f32(t.v[j]);
        s1 += v;
        s2 += v * v;
      }
    }
    block_sum2(s1, s2, red);
    float const mu = s1 * invn;
    float const var = fmaxf(s2 * invn - mu * mu, 0.f);
    float const rs = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      mean[r] = mu;
      rstd[r] = rs;
    }
    T *yr = y + r * N;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int const i = threadIdx.x + p * 256;
      if (i < nvec) {
        Vec16<T> t = *reinterpret_cast<Vec16<T> const *>(src + size_t(i) * 16), o;
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float u = (to_f32(t.v[j]) - mu) * rs;
          o.v[j] = from_f32<T>(u * to_f32(gk[p].v[j]) + to_f32(bk[p].v[j]));
        }
        st_stream(yr + i * VN, o);
      }
    }
    __syncthreads();  // the stage has been read for the last time
    if (threadIdx.x == 0) {
      long long rn = r + (long long)STAGES * gridDim.x;
      if (rn < rows) {
        ptx::fence_proxy_async();
        ptx::mbar_arrive_expect_tx(&bars[st], row_bytes);
        ptx::bulk_load(ring_smem + size_t(st) * row_bytes, x + rn * N, row_bytes, &bars[st]);
      }
    }
    if (++st == STAGES) {
      st = 0;
      phase ^= 1;
    }
  }
}

// Backward for rows of up to 32 columns per lane (N <= 1024 for 16-bit, fp32 alike), one warp per
// row, fed through shared memory: lane 0 of every warp keeps `STAGES` rows of dY, X and the old
// dX in flight with 1-D bulk copies (cp.async.bulk, completion counted on an mbarrier per stage),
// so memory latency is hidden by the ring and not by occupancy — which is what lets each lane keep
// its 64 fp32 dgamma / dbeta accumulators in registers. Each WARP writes its own partial slab.
template <typename T, int PW>
__global__ void __launch_bounds__(128)
    layer_norm_bwd_ring_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                               T *__restrict__ dx, T const *__restrict__ gamma,
                               float const *__restrict__ mean, float const *__restrict__ rstd,
                               float *__restrict__ part, long long rows, int nvec,
                               long long num_warps, int stages, bool dx_acc) {
  constexpr int VN = Vec16<T>::N;
  extern __shared__ __align__(128) uint8_t ring_smem[];
  int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long const N = (long long)nvec * VN;
  uint32_t const row_bytes = uint32_t(N * sizeof(T));
  int const tensors = (dx && dx_acc) ? 3 : 2;  // the old dX travels through the ring only when it is added to
  long long const gw = (long long)blockIdx.x * 4 + warp;
  // layout: [4 warps][stages][tensors][row_bytes] then [4 warps][stages] barriers
  uint8_t *my_ring = ring_smem + size_t(warp) * stages * tensors * row_bytes;
  uint64_t *bars = reinterpret_cast<uint64_t *>(ring_smem + size_t(4) * stages * tensors * row_bytes) +
                   warp * stages;
  if (lane == 0) {
    for (int s2 = 0; s2 < stages; ++s2) ptx::mbar_init(&bars[s2], 1);
    ptx::fence_mbar_init();
  }
  __syncwarp();
  if (gw >= num_warps) return;
  auto issue = [&](long long r, int st) {  // lane 0 only
    uint8_t *dst = my_ring + size_t(st) * tensors * row_bytes;
    ptx::mbar_arrive_expect_tx(&bars[st], tensors * row_bytes);
    ptx::bulk_load(dst, dy + r * N, row_bytes, &bars[st]);
    ptx::bulk_load(dst + row_bytes, x + r * N, row_bytes, &bars[st]);
    if (tensors == 3) ptx::bulk_load(dst + 2 * row_bytes, dx + r * N, row_bytes, &bars[st]);
  };
  if (lane == 0) {
    for (int s2 = 0; s2 < stages; ++s2) {
      long long r = gw + (long long)s2 * num_warps;
      if (r < rows) issue(r, s2);
    }
  }
  float const invn = 1.f / float(N);
  Vec16<T> gk[PW];
  float pg[PW][VN], pb[PW][VN];
#pragma unroll
  for (int p = 0; p < PW; ++p) {
    int i = lane + p * 32;
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      gk[p].v[j] = from_f32<T>(1.f);
      pg[p][j] = 0.f;
      pb[p][j] = 0.f;
    }
    if (gamma && i < nvec) gk[p] = ld_vec(gamma + i * VN);
  }
  int st = 0;
  uint32_t phase = 0;
  for (long long r = gw; r < rows; r += num_warps) {
    ptx::mbar_wait(&bars[st], phase);
    uint8_t const *src = my_ring + size_t(st) * tensors * row_bytes;
    Vec16<T> a[PW], b[PW], o[PW];
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        a[p] = *reinterpret_cast<Vec16<T> const *>(src + size_t(i) * 16);
        b[p] = *reinterpret_cast<Vec16<T> const *>(src + row_bytes + size_t(i) * 16);
        o[p] = Vec16<T>{};
        if (tensors == 3) o[p] = *reinterpret_cast<Vec16<T> const *>(src + 2 * row_bytes + size_t(i) * 16);
      }
    }
    // the row is in registers: hand the stage back and refill it
    __syncwarp();
    if (lane == 0) {
      long long rn = r + (long long)stages * num_warps;
      if (rn < rows) {
        ptx::fence_proxy_async();
        issue(rn, st);
      }
    }
    float const mu = mean[r], rs = rstd[r];
    float ds = 0.f, db = 0.f;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      if (lane + p * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float const dg = to_f32(a[p].v[j]) * to_f32(gk[p].v[j]);
          ds += dg * to_f32(b[p].v[j]);
          db += dg;
        }
      }
    }
    ds = warp_sum(ds);
    db = warp_sum(db);
    float const c1 = (db * mu - ds) * rs * rs * rs * invn;
    float const c2 = -(c1 * mu + db * rs * invn);
    T *dxr = dx ? dx + r * N : nullptr;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float const d = to_f32(a[p].v[j]), xx = to_f32(b[p].v[j]);
          if (dx)
            o[p].v[j] = from_f32<T>(to_f32(o[p].v[j]) + rs * d * to_f32(gk[p].v[j]) + c1 * xx + c2);
          pg[p][j] += d * (xx - mu) * rs;
          pb[p][j] += d;
        }
        if (dx) st_vec(dxr + i * VN, o[p]);
      }
    }
    if (++st == stages) {
      st = 0;
      phase ^= 1;
    }
  }
  if (part) {
    float *pgr = part + (size_t(gw) * 2 + 0) * N;
    float *pbr = part + (size_t(gw) * 2 + 1) * N;
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
#pragma unroll
        for (int q = 0; q < VN / 4; ++q) {
          *reinterpret_cast<float4 *>(pgr + i * VN + q * 4) =
              make_float4(pg[p][q * 4], pg[p][q * 4 + 1], pg[p][q * 4 + 2], pg[p][q * 4 + 3]);
          *reinterpret_cast<float4 *>(pbr + i * VN + q * 4) =
              make_float4(pb[p][q * 4], pb[p][q * 4 + 1], pb[p][q * 4 + 2], pb[p][q * 4 + 3]);
        }
      }
    }
  }
}

// The same ring for LONG rows (up to 16 columns per thread of a 256-thread CTA: N <= 4096): one
// CTA per row at a time, two rows of dY, X and the old dX in flight per CTA as bulk copies, gamma
// in shared memory, the running dgamma / dbeta sums of each thread's columns in registers
// (the older kernel kept them in shared memory and had to rely on occupancy to hide its loads).
template <typename T, int P>
__global__ void __launch_bounds__(256, 2)
    layer_norm_bwd_cta_ring_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                                   T *__restrict__ dx, T const *__restrict__ gamma,
                                   float const *__restrict__ mean, float const *__restrict__ rstd,
                                   float *__restrict__ part, long long rows, int nvec, bool dx_acc) {
  constexpr int VN = Vec16<T>::N, STAGES = 2;
  extern __shared__ __align__(128) uint8_t ring_smem[];
  __shared__ float red[64];
  __shared__ __align__(8) uint64_t bars[STAGES];
  long long const N = (long long)nvec * VN;
  uint32_t const row_bytes = uint32_t(N * sizeof(T));
  int const tensors = (dx && dx_acc) ? 3 : 2;  // the old dX travels through the ring only when it is added to
  T *sgamma = reinterpret_cast<T *>(ring_smem + size_t(STAGES) * tensors * row_bytes);
  if (threadIdx.x == 0) {
    for (int s2 = 0; s2 < STAGES; ++s2) ptx::mbar_init(&bars[s2], 1);
    ptx::fence_mbar_init();
  }
  for (int i = threadIdx.x; i < nvec; i += 256) {
    Vec16<T> g;
#pragma unroll
    for (int j = 0; j < VN; ++j) g.v[j] = from_f32<T>(1.f);
    if (gamma) g = ld_vec(gamma + i * VN);
    *reinterpret_cast<Vec16<T> *>(sgamma + i * VN) = g;
  }
  __syncthreads();
  auto issue = [&](long long r, int st) {  // thread 0 only
    uint8_t *dst = ring_smem + size_t(st) * tensors * row_bytes;
    ptx::mbar_arrive_expect_tx(&bars[st], tensors * row_bytes);
    ptx::bulk_load(dst, dy + r * N, row_bytes, &bars[st]);
    ptx::bulk_load(dst + row_bytes, x + r * N, row_bytes, &bars[st]);
    if (tensors == 3) ptx::bulk_load(dst + 2 * row_bytes, dx + r * N, row_bytes, &bars[st]);
  };
  if (threadIdx.x == 0) {
    for (int s2 = 0; s2 < STAGES; ++s2) {
      long long r = blockIdx.x + (long long)s2 * gridDim.x;
      if (r < rows) issue(r, s2);
    }
  }
  float const invn = 1.f / float(N);
  float pg[P][VN], pb[P][VN];
#pragma unroll
  for (int p = 0; p < P; ++p)
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      pg[p][j] = 0.f;
      pb[p][j] = 0.f;
    }
  int st = 0;
  uint32_t phase = 0;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    ptx::mbar_wait(&bars[st], phase);
    uint8_t const *src = ring_smem + size_t(st) * tensors * row_bytes;
    Vec16<T> a[P], b[P], o[P];
    float ds = 0.f, db = 0.f;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * 256;
      if (i < nvec) {
        a[p] = *reinterpret_cast<Vec16<T> const *>(src + size_t(i) * 16);
        b[p] = *reinterpret_cast<Vec16<T> const *>(src + row_bytes + size_t(i) * 16);
        o[p] = Vec16<T>{};
        if (tensors == 3) o[p] = *reinterpret_cast<Vec16<T> const *>(src + 2 * row_bytes + size_t(i) * 16);
        Vec16<T> g = *reinterpret_cast<Vec16<T> const *>(sgamma + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float const dg = to_f32(a[p].v[j]) * to_f32(g.v[j]);
          ds += dg * to_f32(b[p].v[j]);
          db += dg;
        }
      }
    }
    block_sum2(ds, db, red);  // its barriers also mean: every thread has the row in registers
    if (threadIdx.x == 0) {
      long long rn = r + (long long)STAGES * gridDim.x;
      if (rn < rows) {
        ptx::fence_proxy_async();
        issue(rn, st);
      }
    }
    float const mu = mean[r], rs = rstd[r];
    float const c1 = (db * mu - ds) * rs * rs * rs * invn;
    float const c2 = -(c1 * mu + db * rs * invn);
    T *dxr = dx ? dx + r * N : nullptr;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * 256;
      if (i < nvec) {
        Vec16<T> g = *reinterpret_cast<Vec16<T> const *>(sgamma + i * VN);
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          float const d = to_f32(a[p].v[j]), xx = to_f32(b[p].v[j]);
          if (dx)
            o[p].v[j] = from_f32<T>(to_f32(o[p].v[j]) + rs * d * to_f32(g.v[j]) + c1 * xx + c2);
          pg[p][j] += d * (xx - mu) * rs;
          pb[p][j] += d;
        }
        if (dx) st_vec(dxr + i * VN, o[p]);
      }
    }
    if (++st == STAGES) {
      st = 0;
      phase ^= 1;
    }
  }
  if (part) {
    float *pgr = part + (size_t(blockIdx.x) * 2 + 0) * N;
    float *pbr = part + (size_t(blockIdx.x) * 2 + 1) * N;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      int i = threadIdx.x + p * 256;
      if (i < nvec) {
#pragma unroll
        for (int q = 0; q < VN / 4; ++q) {
          *reinterpret_cast<float4 *>(pgr + i * VN + q * 4) =
              make_float4(pg[p][q * 4], pg[p][q * 4 + 1], pg[p][q * 4 + 2], pg[p][q * 4 + 3]);
          *reinterpret_cast<float4 *>(pbr + i * VN + q * 4) =
              make_float4(pb[p][q * 4], pb[p][q * 4 + 1], pb[p][q * 4 + 2], pb[p][q * 4 + 3]);
        }
      }
    }
  }
}

// 32 columns per lane are 8 packets of fp32 but only 4 of a 16-bit type.
template <typename T> auto widest_ring_kernel() {
  if constexpr (sizeof(T) == 4) return layer_norm_bwd_ring_kernel<T, 8>;
  else return layer_norm_bwd_ring_kernel<T, 4>;
}

template <typename T> auto widest_cta_ring_kernel() {
  if constexpr (sizeof(T) == 4) return layer_norm_bwd_cta_ring_kernel<T, 4>;
  else return layer_norm_bwd_cta_ring_kernel<T, 2>;
}

// Fused backward for short rows, one warp per row. Each lane owns fixed columns, keeps gamma
// and its running dgamma / dbeta sums in registers over all the rows its warp visits; the 8
// warps of a CTA are combined through shared memory into this CTA's partial slab.
template <typename T, int PW>
__global__ void __launch_bounds__(kThreads)
    layer_norm_bwd_warp_rows_kernel(T const *__restrict__ dy, T const *__restrict__ x,
                                    T *__restrict__ dx, T const *__restrict__ gamma,
                                    float const *__restrict__ mean, float const *__restrict__ rstd,
                                    float *__restrict__ part, long long rows, int nvec, bool dx_acc) {
  constexpr int VN = Vec16<T>::N;
  extern __shared__ float4 ln_smem[];  // [2][N] floats, accumulated warp after warp
  int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long const N = (long long)nvec * VN;
  long long const warp0 = (long long)blockIdx.x * (kThreads / 32) + warp;
  long long const wstride = (long long)gridDim.x * (kThreads / 32);
  float const invn = 1.f / float(N);
  float g[PW][VN], pg[PW][VN], pb[PW][VN];
#pragma unroll
  for (int p = 0; p < PW; ++p) {
    int i = lane + p * 32;
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      g[p][j] = 1.f;
      pg[p][j] = 0.f;
      pb[p][j] = 0.f;
    }
    if (gamma && i < nvec) {
      Vec16<T> t = ld_vec(gamma + i * VN);
#pragma unroll
      for (int j = 0; j < VN; ++j) g[p][j] = to_f32(t.v[j]);
    }
  }
  for (long long r = warp0; r < rows; r += wstride) {
    T const *dyr = dy + r * N;
    T const *xr = x + r * N;
    float dv[PW][VN], xv[PW][VN];
    float ds = 0.f, db = 0.f;
    Vec16<T> ak[PW], bk[PW], ok[PW];  // all loads of the row (old dX included) go out together
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32, ii = i < nvec ? i : 0;
      ak[p] = ld_stream(dyr + ii * VN);
      bk[p] = ld_stream(xr + ii * VN);
      ok[p] = Vec16<T>{};
      if (dx && dx_acc) ok[p] = ld_vec(dx + r * N + ii * VN);
    }
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        Vec16<T> const &a = ak[p], &b = bk[p];
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          dv[p][j] = to_f32(a.v[j]);
          xv[p][j] = to_f32(b.v[j]);
          ds += dv[p][j] * xv[p][j] * g[p][j];
          db += dv[p][j] * g[p][j];
        }
      }
    }
    ds = warp_sum(ds);
    db = warp_sum(db);
    float const mu = mean[r], rs = rstd[r];
    float const c1 = (db * mu - ds) * rs * rs * rs * invn;
    float const c2 = -(c1 * mu + db * rs * invn);
#pragma unroll
    for (int p = 0; p < PW; ++p) {
      int i = lane + p * 32;
      if (i < nvec) {
        if (dx) {
          T *dxr = dx + r * N;
          Vec16<T> o = ok[p];
#pragma unroll
          for (int j = 0; j < VN; ++j)
            o.v[j] = from_f32<T>(to_f32(o.v[j]) + rs * dv[p][j] * g[p][j] + c1 * xv[p][j] + c2);
          st_vec(dxr + i * VN, o);
        }
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          pg[p][j] += dv[p][j] * (xv[p][j] - mu) * rs;
          pb[p][j] += dv[p][j];
        }
      }
    }
  }
  if (part) {
    float *sg = reinterpret_cast<float *>(ln_smem);
    float *sb = sg + N;
    for (int w = 0; w < kThreads / 32; ++w) {  // fixed warp order => deterministic
      if (warp == w) {
#pragma unroll
        for (int p = 0; p < PW; ++p) {
          int i = lane + p * 32;
          if (i < nvec) {
#pragma unroll
            for (int j = 0; j < VN; ++j) {
              int c = i * VN + j;
              sg[c] = (w == 0 ? 0.f : sg[c]) + pg[p][j];
              sb[c] = (w == 0 ? 0.f : sb[c]) + pb[p][j];
            }
          }
        }
      }
      __syncthreads();
    }
    float *pgr = part + (size_t(blockIdx.x) * 2 + 0) * N;
    float *pbr = part + (size_t(blockIdx.x) * 2 + 1) * N;
    for (long long c = threadIdx.x; c < N; c += blockDim.x) {
      pgr[c] = sg[c];
      pbr[c] = sb[c];
    }
  }
}

// ================================================================== batch norm (NCHW, fp32)
// One CTA per channel. The channel's data (n * h*w floats, strided by c*h*w between
// samples) is walked three times; after the first walk it is L2 resident.
__global__ void __launch_bounds__(kThreads)
    batch_norm_fwd_kernel(float const *__restrict__ x, float *__restrict__ y,
                          float const *__restrict__ scale, float const *__restrict__ bias,
                          float *__restrict__ run_mean, float *__restrict__ run_var,
                          float *__restrict__ save_mean, float *__restrict__ save_invstd, int n,
                          int c, long long hw, float eps, float factor) {
  __shared__ float red[32];
  int const ch = blockIdx.x;
  long long const m = (long long)n * hw;
  float s = 0.f;
  for (long long i = threadIdx.x; i < m; i += blockDim.x) {
    long long b = i / hw, p = i - b * hw;
    s += x[(b * c + ch) * hw + p];
  }
  float const mu = block_sum(s, red) / float(m);
  float q = 0.f;
  for (long long i = threadIdx.x; i < m; i += blockDim.x) {
    long long b = i / hw, p = i - b * hw;
    float d = x[(b * c + ch) * hw + p] - mu;
    q += d * d;
  }
  float const ssq = block_sum(q, red);
  float const var = ssq / float(m);
  float const invstd = rsqrtf(var + eps);
  if (threadIdx.x == 0) {
    save_mean[ch] = mu;
    save_invstd[ch] = invstd;
    float const unbiased = m > 1 ? ssq / float(m - 1) : var;
    if (run_mean) run_mean[ch] = factor * mu + (1.f - factor) * run_mean[ch];
    if (run_var) run_var[ch] = factor * unbiased + (1.f - factor) * run_var[ch];
  }
  float const g = scale[ch], be = bias[ch];
  for (long long i = threadIdx.x; i < m; i += blockDim.x) {
    long long b = i / hw, p = i - b * hw;
    long long idx = (b * c + ch) * hw + p;
    y[idx] = (x[idx] - mu) * invstd * g + be;
  }
}

__global__ void __launch_bounds__(kThreads)
    batch_norm_bwd_kernel(float const *__restrict__ x, float *__restrict__ dy,
                          float const *__restrict__ y, float *__restrict__ dx,
                          float const *__restrict__ scale, float *__restrict__ dscale,
                          float *__restrict__ dbias, float const *__restrict__ save_mean,
                          float const *__restrict__ save_invstd, int n, int c, long long hw,
                          int relu) {
  __shared__ float red[64];
  int const ch = blockIdx.x;
  long long const m = (long long)n * hw;
  float const mu = save_mean[ch], invstd = save_invstd[ch];
  float sdy = 0.f, sdyx = 0.f;
  for (long long i = threadIdx.x; i < m; i += blockDim.x) {
    long long b = i / hw, p = i - b * hw;
    long long idx = (b * c + ch) * hw + p;
    float d = dy[idx];
    if (relu) {
      d = y[idx] > 0.f ? d : 0.f;
      dy[idx] = d;
    }
    sdy += d;
    sdyx += d * (x[idx] - mu) * invstd;
  }
  block_sum2(sdy, sdyx, red);
  float const g = scale[ch];
  if (threadIdx.x == 0) {
    if (dscale) dscale[ch] += sdyx;
    if (dbias) dbias[ch] += sdy;
  }
  if (dx) {
    float const k = g * invstd / float(m);
    for (long long i = threadIdx.x; i < m; i += blockDim.x) {
      long long b = i / hw, p = i - b * hw;
      long long idx = (b * c + ch) * hw + p;
      float xh = (x[idx] - mu) * invstd;
      dx[idx] += k * (float(m) * dy[idx] - sdy - xh * sdyx);
    }
  }
}

unsigned rows_grid(long long rows, int per_sm) {
  long long cap = (long long)device_info().sm_count * per_sm;
  return (unsigned)std::max<long long>(1, std::min(rows, cap));
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_softmax_forward(void *stream, void const *input, void *output, long long outer,
                        long long dim, long long inner, int dtype) {
  FFK_REQUIRE(outer >= 0 && dim >= 0 && inner >= 0, "softmax: negative extent");
  if (outer == 0 || dim == 0 || inner == 0) return FFK_OK;
  FFK_REQUIRE(input && output, "softmax: null tensor");
  cudaStream_t s = as_stream(stream);
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    bool vec = inner == 1 && dim % VN == 0 && aligned16(input) && aligned16(output) &&
               pick_p(dim / VN) != 0 && dim / VN >= 64;
    bool warp_rows = inner == 1 && dim % VN == 0 && aligned16(input) && aligned16(output) &&
                     pick_pw(dim / VN) != 0;
    if (warp_rows) {
      int nvec = int(dim / VN), p = pick_pw(nvec);
      unsigned grid = rows_grid((outer + 7) / 8, 8);
      FFK_DISPATCH_P(p, P, {
        softmax_warp_rows_kernel<T, P><<<grid, kThreads, 0, s>>>(
            reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), outer, nvec);
      });
    } else if (vec) {
      int nvec = int(dim / VN), p = pick_p(nvec);
      FFK_DISPATCH_P(p, P, {
        softmax_rows_kernel<T, P><<<rows_grid(outer, 8), kThreads, 0, s>>>(
            reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), outer, nvec);
      });
    } else if (inner == 1 && dim >= 1024 && size_t(dim) * sizeof(T) <= (96u << 10)) {
      size_t smem = (size_t(dim) * sizeof(T) + 15) / 16 * 16;
      auto kern = softmax_rows_smem_kernel<T>;
      FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int per_sm = 0;
      FFK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem));
      kern<<<rows_grid(outer, std::max(1, per_sm)), kThreads, smem, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), outer, dim);
    } else if (inner >= 32) {
      unsigned grid = stream_grid(size_t(outer * inner), kThreads, 8);
      softmax_strided_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), outer, dim, inner);
    } else {
      unsigned grid = stream_grid(size_t(outer * inner) * 32, kThreads, 8);
      softmax_lines_kernel<T><<<grid, kThreads, 0, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output), outer, dim, inner);
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_softmax_backward(void *stream, void *input_grad, void const *output_grad,
                         void const *output, long long outer, long long dim, long long inner,
                         int dtype, int mode) {
  FFK_REQUIRE(outer >= 0 && dim >= 0 && inner >= 0, "softmax: negative extent");
  FFK_REQUIRE(mode == 0 || mode == 1, "softmax backward: mode must be 0 or 1");
  size_t n = size_t(outer) * size_t(dim) * size_t(inner);
  if (n == 0) return FFK_OK;
  FFK_REQUIRE(input_grad && output_grad, "softmax backward: null tensor");
  if (mode == 0) return ffk_copy(stream, input_grad, output_grad, n, dtype);
  FFK_REQUIRE(output, "softmax backward: Jacobian mode needs `output`");
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    bool warp_rows = inner == 1 && dim % VN == 0 && aligned16(input_grad) && aligned16(output_grad) &&
                     aligned16(output) && pick_pw(dim / VN) != 0;
    if (warp_rows) {
      int nvec = int(dim / VN), p = pick_pw(nvec);
      unsigned grid = rows_grid((outer + 7) / 8, 8);
      FFK_DISPATCH_P(p, P, {
        softmax_jacobian_warp_rows_kernel<T, P><<<grid, kThreads, 0, as_stream(stream)>>>(
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(output_grad),
            reinterpret_cast<T const *>(output), outer, nvec);
      });
    } else {
      unsigned grid = stream_grid(size_t(outer * inner) * 32, kThreads, 8);
      softmax_jacobian_kernel<T><<<grid, kThreads, 0, as_stream(stream)>>>(
          reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(output_grad),
          reinterpret_cast<T const *>(output), outer, dim, inner);
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_layer_norm_forward(void *stream, void const *input, void *output, void const *gamma,
                           void const *beta, float *mean, float *rstd, long long M, long long N,
                           float eps, int dtype) {
  FFK_REQUIRE(M >= 0 && N >= 0, "layer_norm: negative extent");
  if (M == 0 || N == 0) return FFK_OK;
  FFK_REQUIRE(input && output && mean && rstd, "layer_norm: null tensor");
  cudaStream_t s = as_stream(stream);
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    bool vec = N % VN == 0 && aligned16(input) && aligned16(output) &&
               (!gamma || aligned16(gamma)) && (!beta || aligned16(beta)) && pick_p(N / VN) != 0;
    if (vec && pick_pw(N / VN) != 0 && pick_pw(N / VN) <= 4) {
      int nvec = int(N / VN), p = pick_pw(nvec);
      unsigned grid = rows_grid((M + 7) / 8, 8);
      FFK_DISPATCH_P(p, P, {
        layer_norm_fwd_warp_rows_kernel<T, P><<<grid, kThreads, 0, s>>>(
            reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
            reinterpret_cast<T const *>(gamma), reinterpret_cast<T const *>(beta), mean, rstd, M,
            nvec, eps);
      });
    } else if (vec && std::getenv("FFK_LN_NO_RING") == nullptr && N * sizeof(T) >= 8192 &&
               N * sizeof(T) <= 32768) {
      int nvec = int(N / VN);
      size_t smem = size_t(3) * size_t(N) * sizeof(T);
      auto launch = [&](auto kern) -> int {
        FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        FFK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem));
        kern<<<rows_grid(M, std::max(1, per_sm)), 256, smem, s>>>(
            reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
            reinterpret_cast<T const *>(gamma), reinterpret_cast<T const *>(beta), mean, rstd, M, nvec, eps);
        return FFK_OK;
      };
      int const packets_per_thread = (nvec + 255) / 256;   // rows of 8 .. 32 KB: 2 .. 8 packets of 16 bytes
      int rc = packets_per_thread <= 2   ? launch(layer_norm_fwd_cta_ring_kernel<T, 2>)
               : packets_per_thread <= 4 ? launch(layer_norm_fwd_cta_ring_kernel<T, 4>)
                                         : launch(layer_norm_fwd_cta_ring_kernel<T, 8>);
      if (rc) return rc;
    } else if (vec) {
      int nvec = int(N / VN), p = pick_p(nvec);
      FFK_DISPATCH_P(p, P, {
        layer_norm_fwd_rows_kernel<T, P><<<rows_grid(M, 6), kThreads, 0, s>>>(
            reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
            reinterpret_cast<T const *>(gamma), reinterpret_cast<T const *>(beta), mean, rstd, M,
            nvec, eps);
      });
    } else {
      layer_norm_fwd_generic_kernel<T><<<rows_grid(M, 8), kThreads, 0, s>>>(
          reinterpret_cast<T const *>(input), reinterpret_cast<T *>(output),
          reinterpret_cast<T const *>(gamma), reinterpret_cast<T const *>(beta), mean, rstd, M, N,
          eps);
    }
  });
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

size_t ffk_layer_norm_backward_workspace_bytes(long long N) {
  if (N <= 0) return 0;
  return size_t(device_info().sm_count) * 12 /*CTAs per SM*/ * 2 /*dgamma,dbeta*/ * size_t(N) *
         sizeof(float);
}

}  // extern "C"

namespace {
// dx_acc: input_grad += ... (the reference's beta = 1 convention) or input_grad = ... (first
// writer of that gradient: no zero fill before, no read of the old value here)
int layer_norm_backward_impl(void *stream, void const *output_grad, void const *input,
                             void *input_grad, void const *gamma, void *gamma_grad,
                             void *beta_grad, float const *mean, float const *rstd, long long M,
                             long long N, int dtype, int accumulate_param_grads, void *workspace,
                             size_t workspace_bytes, bool dx_acc) {
  FFK_REQUIRE(M >= 0 && N >= 0, "layer_norm: negative extent");
  if (M == 0 || N == 0) return FFK_OK;
  FFK_REQUIRE(output_grad && input && mean && rstd, "layer_norm backward: null tensor");
  cudaStream_t s = as_stream(stream);
  bool want_params = gamma_grad || beta_grad;
  FFK_DISPATCH_FLOAT(dtype, T, {
    constexpr int VN = Vec16<T>::N;
    bool vec = N % VN == 0 && aligned16(output_grad) && aligned16(input) &&
               (!input_grad || aligned16(input_grad)) && (!gamma || aligned16(gamma)) &&
               pick_p(N / VN) != 0;
    size_t per_part = size_t(2) * size_t(N) * sizeof(float);
    // partial slabs are stored with float4: a workspace that is not 16-byte aligned is not used
    size_t parts_fit = (workspace && want_params && aligned16(workspace)) ? workspace_bytes / per_part : 0;
    bool fused = vec && want_params && parts_fit >= 1;
    bool short_rows = vec && pick_pw(N / VN) != 0 && pick_pw(N / VN) <= 2;
    static bool const no_ring = std::getenv("FFK_LN_NO_RING") != nullptr;
    int const ring_pw = vec ? pick_pw(N / VN) : 0;
    // measured: the ring wins from 256-byte rows up for 16-bit types and from 1 KB rows for fp32
    bool ring_rows = !no_ring && ring_pw != 0 && ring_pw * VN <= 32 &&
                     (N * sizeof(T)) >= (sizeof(T) == 2 ? 256u : 1024u);
    if (ring_rows && (input_grad || fused)) {
      int nvec = int(N / VN);
      size_t row_bytes = size_t(N) * sizeof(T);
      int tensors = (input_grad && dx_acc) ? 3 : 2;
      int stages = (int)std::max<size_t>(2, std::min<size_t>(6, (24 * 1024) / (tensors * row_bytes)));
      size_t smem = size_t(4) * stages * tensors * row_bytes + size_t(4) * stages * 8;
      float *wpart = fused ? reinterpret_cast<float *>(workspace) : nullptr;
      long long warps = 0;
      auto launch = [&](auto kern) -> int {
        FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;  // registers (up to 240 a thread at 32 columns per lane) or the ring decide
        FFK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
        warps = (long long)device_info().sm_count * std::max(1, std::min(per_sm, 4)) * 4;
        warps = std::min<long long>(warps, M);
        if (fused) warps = std::min<long long>(warps, (long long)parts_fit);
        kern<<<unsigned((warps + 3) / 4), 128, smem, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, wpart, M,
            nvec, warps, stages, dx_acc);
        return FFK_OK;
      };
      int rc = FFK_OK;
      switch (ring_pw) {
        case 1: rc = launch(layer_norm_bwd_ring_kernel<T, 1>); break;
        case 2: rc = launch(layer_norm_bwd_ring_kernel<T, 2>); break;
        case 4: rc = launch(layer_norm_bwd_ring_kernel<T, 4>); break;
        default: rc = launch(widest_ring_kernel<T>()); break;  // reached by fp32 only (8 x 4 columns)
      }
      if (rc) return rc;
      FFK_LAUNCH_CHECK();
      if (fused) {
        layer_norm_param_reduce_kernel<T><<<unsigned((N + 31) / 32), 1024, 0, s>>>(
            reinterpret_cast<float const *>(workspace), int(warps), N,
            reinterpret_cast<T *>(gamma_grad), reinterpret_cast<T *>(beta_grad),
            accumulate_param_grads != 0);
        FFK_LAUNCH_CHECK();
      }
    } else if (!no_ring && vec && pick_p(N / VN) != 0 && pick_p(N / VN) * VN <= 16 &&
               N * sizeof(T) >= 4096 && (input_grad || fused)) {
      // long rows: CTA-level ring (N <= 4096 at 16 columns per thread)
      int nvec = int(N / VN), pp = pick_p(nvec);
      size_t row_bytes = size_t(N) * sizeof(T);
      int tensors = (input_grad && dx_acc) ? 3 : 2;
      size_t smem = size_t(2) * tensors * row_bytes + row_bytes;
      float *wpart = fused ? reinterpret_cast<float *>(workspace) : nullptr;
      unsigned grid = 0;
      auto launch = [&](auto kern) -> int {
        FFK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        FFK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, smem));
        grid = rows_grid(M, std::max(1, per_sm));
        if (fused) grid = (unsigned)std::min<size_t>(grid, parts_fit);
        kern<<<grid, 256, smem, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, wpart, M, nvec, dx_acc);
        return FFK_OK;
      };
      int rc = FFK_OK;
      switch (pp) {
        case 1: rc = launch(layer_norm_bwd_cta_ring_kernel<T, 1>); break;
        case 2: rc = launch(layer_norm_bwd_cta_ring_kernel<T, 2>); break;
        default: rc = launch(widest_cta_ring_kernel<T>()); break;  // reached by fp32 only
      }
      if (rc) return rc;
      FFK_LAUNCH_CHECK();
      if (fused) {
        layer_norm_param_reduce_kernel<T><<<unsigned((N + 31) / 32), 1024, 0, s>>>(
            reinterpret_cast<float const *>(workspace), int(grid), N,
            reinterpret_cast<T *>(gamma_grad), reinterpret_cast<T *>(beta_grad),
            accumulate_param_grads != 0);
        FFK_LAUNCH_CHECK();
      }
    } else if (short_rows && (input_grad || fused)) {
      int nvec = int(N / VN), p = pick_pw(nvec);
      size_t smem = size_t(2) * size_t(N) * sizeof(float);
      unsigned grid = rows_grid((M + 7) / 8, 3);
      if (fused) grid = (unsigned)std::min<size_t>(grid, parts_fit);
      float *wpart = fused ? reinterpret_cast<float *>(workspace) : nullptr;
      if (p == 1) {
        layer_norm_bwd_warp_rows_kernel<T, 1><<<grid, kThreads, smem, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, wpart, M, nvec, dx_acc);
      } else {
        layer_norm_bwd_warp_rows_kernel<T, 2><<<grid, kThreads, smem, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, wpart, M, nvec, dx_acc);
      }
      FFK_LAUNCH_CHECK();
      if (fused) {
        layer_norm_param_reduce_kernel<T><<<unsigned((N + 31) / 32), 1024, 0, s>>>(
            reinterpret_cast<float const *>(workspace), int(grid), N,
            reinterpret_cast<T *>(gamma_grad), reinterpret_cast<T *>(beta_grad),
            accumulate_param_grads != 0);
        FFK_LAUNCH_CHECK();
      }
    } else if (vec && (input_grad || fused)) {
      int nvec = int(N / VN);
      bool narrow = nvec <= 128;  // rows of <= 128 packets: 128-thread CTAs keep every lane busy
      int p = narrow ? 1 : pick_p(nvec);
      size_t smem = size_t(3) * size_t(N) * sizeof(float);
      int per_sm = (int)std::max<size_t>(1, std::min<size_t>(narrow ? 12 : 4, (200 * 1024) / (smem + 1024)));
      unsigned grid = rows_grid(M, per_sm);
      if (fused) grid = (unsigned)std::min<size_t>(grid, parts_fit);
      float *part = fused ? reinterpret_cast<float *>(workspace) : nullptr;
      if (narrow) {
        FFK_CUDA(cudaFuncSetAttribute(layer_norm_bwd_rows_kernel<T, 1, 128>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        layer_norm_bwd_rows_kernel<T, 1, 128><<<grid, 128, smem, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
            reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, part, M, nvec, dx_acc);
      } else {
        FFK_DISPATCH_P(p, P, {
          FFK_CUDA(cudaFuncSetAttribute(layer_norm_bwd_rows_kernel<T, P, kThreads>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          layer_norm_bwd_rows_kernel<T, P, kThreads><<<grid, kThreads, smem, s>>>(
              reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
              reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, part, M, nvec, dx_acc);
        });
      }
      FFK_LAUNCH_CHECK();
      if (fused) {
        layer_norm_param_reduce_kernel<T><<<unsigned((N + 31) / 32), 1024, 0, s>>>(
            reinterpret_cast<float const *>(workspace), int(grid), N,
            reinterpret_cast<T *>(gamma_grad), reinterpret_cast<T *>(beta_grad),
            accumulate_param_grads != 0);
        FFK_LAUNCH_CHECK();
      }
    } else if (input_grad) {
      layer_norm_bwd_generic_kernel<T><<<rows_grid(M, 8), kThreads, 0, s>>>(
          reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input),
          reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(gamma), mean, rstd, M, N, dx_acc);
      FFK_LAUNCH_CHECK();
    }
    if (want_params && !fused) {
      layer_norm_param_columns_kernel<T><<<unsigned((N + 31) / 32), kThreads, 0, s>>>(
          reinterpret_cast<T const *>(output_grad), reinterpret_cast<T const *>(input), mean, rstd,
          reinterpret_cast<T *>(gamma_grad), reinterpret_cast<T *>(beta_grad), M, N,
          accumulate_param_grads != 0);
      FFK_LAUNCH_CHECK();
    }
  });
  return FFK_OK;
}
}  // namespace

extern "C" {

int ffk_layer_norm_backward(void *stream, void const *output_grad, void const *input,
                            void *input_grad, void const *gamma, void *gamma_grad,
                            void *beta_grad, float const *mean, float const *rstd, long long M,
                            long long N, int dtype, int accumulate_param_grads, void *workspace,
                            size_t workspace_bytes) {
  return layer_norm_backward_impl(stream, output_grad, input, input_grad, gamma, gamma_grad, beta_grad, mean,
                                  rstd, M, N, dtype, accumulate_param_grads, workspace, workspace_bytes, true);
}

int ffk_layer_norm_backward_ex(void *stream, void const *output_grad, void const *input,
                               void *input_grad, void const *gamma, void *gamma_grad,
                               void *beta_grad, float const *mean, float const *rstd, long long M,
                               long long N, int dtype, int accumulate_param_grads,
                               int accumulate_input_grad, void *workspace, size_t workspace_bytes) {
  return layer_norm_backward_impl(stream, output_grad, input, input_grad, gamma, gamma_grad, beta_grad, mean,
                                  rstd, M, N, dtype, accumulate_param_grads, workspace, workspace_bytes,
                                  accumulate_input_grad != 0);
}

int ffk_batch_norm_forward(void *stream, float const *input, float *output, float const *scale,
                           float const *bias, float *running_mean, float *running_var,
                           float *save_mean, float *save_invstd, int n, int c, int h, int w,
                           float eps, float factor) {
  FFK_REQUIRE(n >= 0 && c >= 0 && h >= 0 && w >= 0, "batch_norm: negative extent");
  if (n == 0 || c == 0 || h == 0 || w == 0) return FFK_OK;
  FFK_REQUIRE(input && output && scale && bias && save_mean && save_invstd,
              "batch_norm: null tensor");
  batch_norm_fwd_kernel<<<c, kThreads, 0, as_stream(stream)>>>(
      input, output, scale, bias, running_mean, running_var, save_mean, save_invstd, n, c,
      (long long)h * w, eps, factor);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}

int ffk_batch_norm_backward(void *stream, float const *input, float *output_grad,
                            float const *output, float *input_grad, float const *scale,
                            float *scale_grad, float *bias_grad, float const *save_mean,
                            float const *save_invstd, int n, int c, int h, int w, int relu) {
  FFK_REQUIRE(n >= 0 && c >= 0 && h >= 0 && w >= 0, "batch_norm: negative extent");
  if (n == 0 || c == 0 || h == 0 || w == 0) return FFK_OK;
  FFK_REQUIRE(input && output_grad && scale && save_mean && save_invstd,
              "batch_norm backward: null tensor");
  FFK_REQUIRE(!relu || output, "batch_norm backward: relu needs `output`");
  batch_norm_bwd_kernel<<<c, kThreads, 0, as_stream(stream)>>>(
      input, output_grad, output, input_grad, scale, scale_grad, bias_grad, save_mean,
      save_invstd, n, c, (long long)h * w, relu);
  FFK_LAUNCH_CHECK();
  return FFK_OK;
}


// ------------------------------------------------------------------------------------------
// LayerNorm backward for callers that cannot allocate.
//
// Who that is: code that runs as a task under a runtime which has already taken the GPU's
// memory for its own pool — the reference's legacy Legion/Realm runtime reserves the
// framebuffer when the process starts, so a cudaMalloc issued later from inside a task can
// simply fail, and even when it succeeds it synchronises the whole device and stalls every
// other task's stream — and callers that reach the backward call holding neither an
// Allocator nor a PerDeviceFFHandle workspace to lend. For them the per-CTA partial sums of
// dgamma / dbeta, which ffk_layer_norm_backward keeps in caller scratch, live here, in a
// fixed array that is part of the library image. Every GPU that loads libffk.so carries these
// 4 MiB until the process exits.
//
// The array's NAME is only a host-side handle: handing it to a kernel from host code would
// pass a host address. Its device address is looked up once per device with
// cudaGetSymbolAddress (below) and that pointer is what the kernels receive.
constexpr size_t kStaticScratchFloats = size_t(1) << 20;   // 4 MiB
constexpr long long kStaticMaxColumns = 16384;             // => at least 32 partial slabs
constexpr int kStaticMaxDevices = 64;

namespace {
static __device__ __align__(256) float g_layer_norm_static_scratch[kStaticScratchFloats];
float *g_static_scratch_addr[kStaticMaxDevices];  // device address per device ordinal, lazily filled
std::mutex g_static_scratch_mutex;

// Resolves (once per device, with that device current) the device address of the array.
int static_scratch_of_current_device(float **out) {
  int dev = -1;
  FFK_CUDA(cudaGetDevice(&dev));
  FFK_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "layer_norm static: device ordinal %d unsupported", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_addr[dev] == nullptr) {
    float *addr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&addr), g_layer_norm_static_scratch);
    if (e != cudaSuccess) return ::ffk::cuda_fail(e, "cudaGetSymbolAddress(g_layer_norm_static_scratch)", __FILE__, __LINE__);
    FFK_REQUIRE(addr != nullptr, "layer_norm static: symbol address resolved to null");
    g_static_scratch_addr[dev] = addr;
  }
  *out = g_static_scratch_addr[dev];
  return FFK_OK;
}
}  // namespace

// Same arithmetic and the same kernels as ffk_layer_norm_backward; the only difference is
// where the partial sums are kept. Nothing here (or below it) calls cudaMalloc, cudaFree or
// any allocating library routine.
//  * The first call on a device resolves the array's address (see above). Make that first
//    call outside any stream capture.
//  * Limit: N <= 16384 columns; larger rows are refused with FFK_ERR_LIMIT.
//  * All calls on one GPU share the one array and it keeps whatever the previous call left
//    there. That is harmless for sequential calls — each call's first kernel overwrites every
//    slab that its second kernel reads, so nothing has to be cleared — but two calls whose
//    kernels overlap on the GPU (different streams, no ordering between them) would trample
//    each other's partial sums. Order such calls on one stream or with events.
int ffk_layer_norm_backward_static(void *stream, void const *output_grad, void const *input,
                                   void *input_grad, void const *gamma, void *gamma_grad,
                                   void *beta_grad, float const *mean, float const *rstd,
                                   long long M, long long N, int dtype,
                                   int accumulate_param_grads) {
  if (N > kStaticMaxColumns) {
    set_error("layer_norm static: N = %lld exceeds the %lld columns the built-in 4 MiB scratch "
              "is sized for; use ffk_layer_norm_backward with a caller workspace",
              N, kStaticMaxColumns);
    return FFK_ERR_LIMIT;
  }
  float *scratch = nullptr;
  int rc = static_scratch_of_current_device(&scratch);
  if (rc) return rc;
  return ffk_layer_norm_backward(stream, output_grad, input, input_grad, gamma, gamma_grad,
                                 beta_grad, mean, rstd, M, N, dtype, accumulate_param_grads,
                                 scratch, kStaticScratchFloats * sizeof(float));
}

}  // extern "C"
