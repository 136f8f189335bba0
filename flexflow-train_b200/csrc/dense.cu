// Linear and BatchMatmul on top of the GEMM engine (gemm_sm100.cu).
#include "gemm_sm100.cuh"
#include "philox.cuh"

namespace ffk {
namespace {

int run_gemm(gemm::Problem const &p, gemm::Epilogue const &e, cudaStream_t s, int path) {
  switch (path) {
    case FFK_GEMM_AUTO: return gemm::launch(p, e, s);
    case FFK_GEMM_TENSOR: return gemm::launch_tc(p, e, s);
    case FFK_GEMM_SIMT: return gemm::launch_simt(p, e, s);
    case FFK_GEMM_SPLIT_BF16: return gemm::launch_split_bf16_registered(p, e, s);
    default: set_error("gemm: unknown path %d", path); return FFK_ERR_INVALID;
  }
}

// dY <- dY * act'(.) in place. `ref` is Y for relu/sigmoid/tanh and the
// pre-activation for gelu.
// With `mask` (the keep bits of a dropout that followed the activation, 1 bit per element) the
// dropout gradient is folded in: dY <- dY * keep * scale * act'(.), one pass instead of two.
template <typename T>
__global__ void __launch_bounds__(256)
    dropout_act_backward_kernel(T *__restrict__ dy, T const *__restrict__ ref,
                                uint32_t const *__restrict__ mask, size_t n, int act, float scale) {
  constexpr int VN = Vec16<T>::N;
  size_t nvec = n / VN;
  size_t stride = size_t(gridDim.x) * blockDim.x;
  auto grad = [&](float gv, float yv) -> float {
    if (act == FFK_ACT_NONE) return gv;
    if (act == FFK_ACT_RELU) return yv > 0.f ? gv : 0.f;
    if (act == FFK_ACT_SIGMOID) return gv * yv * (1.f - yv);
    if (act == FFK_ACT_TANH) return gv * (1.f - yv * yv);
    constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
    float u = yv * (C * yv * yv + B);
    float t = tanhf(u);
    float du = B + 3.f * C * yv * yv;
    return gv * (0.5f + 0.5f * t + 0.5f * yv * (1.f - t * t) * du);
  };
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    size_t const e0 = i * VN;
    uint32_t const bits = __ldg(mask + e0 / 32) >> (e0 % 32);
    Vec16<T> g = ld_vec(dy + e0);
    Vec16<T> y;
    if (act != FFK_ACT_NONE) y = ld_stream(ref + e0);
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      float gv = ((bits >> j) & 1u) ? to_f32(g.v[j]) * scale : 0.f;
      g.v[j] = from_f32<T>(grad(gv, act != FFK_ACT_NONE ? to_f32(y.v[j]) : 0.f));
    }
    st_vec(dy + e0, g);
  }
  for (size_t e = nvec * VN + size_t(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
    float gv = ((mask[e / 32] >> (e % 32)) & 1u) ? to_f32(dy[e]) * scale : 0.f;
    dy[e] = from_f32<T>(grad(gv, act != FFK_ACT_NONE ? to_f32(ref[e]) : 0.f));
  }
}

// dx <- gate > 0 ? dx * scale : 0 (the gradient through a ReLU, or ReLU + dropout, whose stored
// OUTPUT is `gate`): the stand-alone form of the data-gradient epilogue's gate mode.
template <typename T>
__global__ void __launch_bounds__(256)
    gate_kernel(T *__restrict__ dx, T const *__restrict__ gate, size_t n, float scale) {
  constexpr int VN = Vec16<T>::N;
  size_t nvec = n / VN, stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    Vec16<T> g = ld_vec(dx + i * VN), y = ld_stream(gate + i * VN);
#pragma unroll
    for (int j = 0; j < VN; ++j) g.v[j] = from_f32<T>(to_f32(y.v[j]) > 0.f ? to_f32(g.v[j]) * scale : 0.f);
    st_vec(dx + i * VN, g);
  }
  for (size_t e = nvec * VN + size_t(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride)
    dx[e] = from_f32<T>(to_f32(gate[e]) > 0.f ? to_f32(dx[e]) * scale : 0.f);
}

template <typename T>
__global__ void __launch_bounds__(256)
    act_backward_kernel(T *__restrict__ dy, T const *__restrict__ ref, size_t n, int act) {
  constexpr int VN = Vec16<T>::N;
  size_t nvec = n / VN;
  size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    Vec16<T> g = ld_vec(dy + i * VN);
    Vec16<T> y = ld_stream(ref + i * VN);
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      float gv = to_f32(g.v[j]), yv = to_f32(y.v[j]), o;
      if (act == FFK_ACT_RELU) o = yv > 0.f ? gv : 0.f;
      else if (act == FFK_ACT_SIGMOID) o = gv * yv * (1.f - yv);
      else if (act == FFK_ACT_TANH) o = gv * (1.f - yv * yv);
      else {
        constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
        float u = yv * (C * yv * yv + B);
        float t = tanhf(u);
        float du = B + 3.f * C * yv * yv;
        o = gv * (0.5f + 0.5f * t + 0.5f * yv * (1.f - t * t) * du);
      }
      g.v[j] = from_f32<T>(o);
    }
    st_vec(dy + i * VN, g);
  }
  // tail
  for (size_t i = nvec * VN + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float gv = to_f32(dy[i]), yv = to_f32(ref[i]), o;
    if (act == FFK_ACT_RELU) o = yv > 0.f ? gv : 0.f;
    else if (act == FFK_ACT_SIGMOID) o = gv * yv * (1.f - yv);
    else if (act == FFK_ACT_TANH) o = gv * (1.f - yv * yv);
    else {
      constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
      float u = yv * (C * yv * yv + B);
      float t = tanhf(u);
      float du = B + 3.f * C * yv * yv;
      o = gv * (0.5f + 0.5f * t + 0.5f * yv * (1.f - t * t) * du);
    }
    dy[i] = from_f32<T>(o);
  }
}

// db[o] += sum_b dY[b][o]. One CTA owns a strip of 32 columns and walks every
// row with 8 row-lanes; fixed summation order => run-to-run deterministic.
template <typename T>
__global__ void __launch_bounds__(256)
    column_sum_accumulate_kernel(T const *__restrict__ dy, T *__restrict__ db, int rows,
                                 int cols) {
  __shared__ float part[8][33];
  int const cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  int const col = blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (col < cols) {
    for (int r = ry; r < rows; r += 8) acc += to_f32(dy[size_t(r) * cols + col]);
  }
  part[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && col < cols) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += part[i][cx];
    db[col] = from_f32<T>(to_f32(db[col]) + s);
  }
}


// Bias gradient at HBM speed: phase 1 sums a slab of rows per CTA into fp32 partials
// part[split][col] (16-byte loads, 32 lanes x 8 row-lanes), phase 2 adds the partials to db
// in split order. Deterministic; needs split*cols floats of scratch.
template <typename T>
__global__ void __launch_bounds__(256)
    column_sum_partial_kernel(T const *__restrict__ dy, float *__restrict__ part, int rows,
                              int cols, int rows_per_split) {
  constexpr int VN = Vec16<T>::N;
  __shared__ float sm[8][32 * VN + 1];
  int const cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  int const col = (blockIdx.x * 32 + cx) * VN;
  int const r0 = blockIdx.y * rows_per_split;
  int const r1 = min(rows, r0 + rows_per_split);
  float acc[VN];
#pragma unroll
  for (int j = 0; j < VN; ++j) acc[j] = 0.f;
  if (col < cols) {
    for (int r = r0 + ry; r < r1; r += 8) {
      Vec16<T> v = ld_stream(dy + size_t(r) * cols + col);
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] += to_f32(v.v[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < VN; ++j) sm[ry][cx * VN + j] = acc[j];
  __syncthreads();
  for (int c = threadIdx.x; c < 32 * VN; c += 256) {
    int gc = blockIdx.x * 32 * VN + c;
    if (gc < cols) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t += sm[i][c];
      part[size_t(blockIdx.y) * cols + gc] = t;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    column_sum_finish_kernel(float const *__restrict__ part, T *__restrict__ db, int cols,
                             int splits) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float t = 0.f;
  for (int s = 0; s < splits; ++s) t += part[size_t(s) * cols + c];
  db[c] = from_f32<T>(to_f32(db[c]) + t);
}

// Row splits of the bias-gradient reduction: enough (column block x split) CTAs to cover the
// machine about four times, whatever the width; 32 for wide layers, up to 256 for narrow ones.
inline int bias_grad_max_splits(int out_dim) {
  int col_blocks = (out_dim + 255) / 256;  // 32 lanes x 8 sixteen-bit columns per CTA
  int want = (4 * device_info().sm_count + col_blocks - 1) / col_blocks;
  return std::max(32, std::min(256, want));
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

int ffk_set_gemm_workspace(void *workspace, size_t workspace_bytes) {
  FFK_REQUIRE(workspace == nullptr || aligned16(workspace), "gemm workspace must be 16-byte aligned");
  return gemm::set_workspace(workspace, workspace ? workspace_bytes : 0);
}

int ffk_set_gemm_fp32_tensor_math(int allow) { return gemm::set_fp32_tensor_math(allow); }

int ffk_set_gemm_kcut(int enable) { return gemm::set_kcut(enable); }

int ffk_gemm_kcut_plan(int M, int N, int K, int pairs, int *plan5) {
  return gemm::kcut_plan(M, N, K, pairs, plan5);
}
int ffk_gemm_kcut_segments(int M, int N, int K, int pairs, int pair, int *six, int max_segments) {
  if (pairs <= 0 || pair < 0 || pair >= pairs || max_segments < 0) {
    set_error("ffk_gemm_kcut_segments: pair %d of %d", pair, pairs);
    return FFK_ERR_INVALID;
  }
  return gemm::kcut_segments(M, N, K, pairs, pair, six, max_segments);
}

int ffk_gemm(void *stream, void const *a, void const *b, void *d, int M, int N, int K, int batch,
             long long lda, long long ldb, long long ldd, long long a_batch_stride,
             long long b_batch_stride, long long d_batch_stride, int in_dtype, int out_dtype,
             int a_mn_major, int b_mn_major, float alpha, int accumulate, void const *bias,
             int activation, int path) {
  FFK_REQUIRE(M >= 0 && N >= 0 && K >= 0 && batch >= 0, "ffk_gemm: negative extent");
  if (M == 0 || N == 0 || batch == 0) return FFK_OK;
  FFK_REQUIRE(a && b && d, "ffk_gemm: null operand");
  gemm::Problem p{a, b, lda, ldb, a_batch_stride, b_batch_stride, M, N, K, batch,
                  in_dtype, out_dtype, a_mn_major != 0, b_mn_major != 0};
  gemm::Epilogue e{d, ldd, d_batch_stride, bias, activation, accumulate, alpha, 0.f, nullptr};
  return run_gemm(p, e, as_stream(stream), path);
}

int ffk_linear_forward(void *stream, void const *input, void *output, void const *weight,
                       void const *bias, int in_dim, int out_dim, int batch_size, int dtype,
                       int activation, int path) {
  FFK_REQUIRE(in_dim > 0 && out_dim > 0 && batch_size >= 0, "linear forward: bad extents");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output && weight, "linear forward: null tensor");
  gemm::Problem p{input, weight, in_dim, in_dim, 0, 0, batch_size, out_dim, in_dim, 1,
                  dtype, dtype, false, false};
  gemm::Epilogue e{output, out_dim, 0, bias, activation, 0, 1.f, 0.f, nullptr};
  return run_gemm(p, e, as_stream(stream), path);
}

}  // extern "C"

namespace {
int linear_backward_impl(void *stream, void const *input, void *input_grad, void const *output,
                         void *output_grad, void const *weight, void *weight_grad,
                         void *bias_grad, int in_dim, int out_dim, int batch_size, int dtype,
                         int activation, float l2_lambda, void const *preact, int path,
                         void *workspace, size_t workspace_bytes,
                         ffk_linear_backward_fusion_t const &fu) {
  FFK_REQUIRE(in_dim > 0 && out_dim > 0 && batch_size >= 0, "linear backward: bad extents");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output_grad && weight && weight_grad, "linear backward: null tensor");
  cudaStream_t s = as_stream(stream);
  size_t n_out = size_t(out_dim) * batch_size;

  if (fu.output_grad_is_preactivation) {
    // a gated data-gradient epilogue upstream already applied this layer's activation (and
    // dropout) gradient: nothing to do here
  } else if (fu.dropout_mask != nullptr && fu.dropout_rate > 0.f) {
    void const *ref = activation == FFK_ACT_GELU ? preact : output;
    if (activation == FFK_ACT_GELU && !preact) {
      set_error("linear backward: GELU gradient needs the pre-activation buffer");
      return FFK_ERR_UNSUPPORTED;
    }
    FFK_REQUIRE(activation == FFK_ACT_NONE || ref, "linear backward: activation gradient needs `output`");
    FFK_REQUIRE(aligned16(output_grad) && (!ref || aligned16(ref)), "linear backward: unaligned tensor");
    FFK_REQUIRE(fu.dropout_rate <= 1.f, "linear backward: dropout rate %f outside [0,1]", fu.dropout_rate);
    float scale = fu.dropout_rate < 1.f ? 1.f / (1.f - fu.dropout_rate) : 0.f;
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = stream_grid(n_out / Vec16<T>::N + 1, 256, 8);
      dropout_act_backward_kernel<T><<<grid, 256, 0, s>>>(
          reinterpret_cast<T *>(output_grad), reinterpret_cast<T const *>(ref),
          reinterpret_cast<uint32_t const *>(fu.dropout_mask), n_out, activation, scale);
    });
    FFK_LAUNCH_CHECK();
  } else if (activation != FFK_ACT_NONE) {
    void const *ref = output;
    if (activation == FFK_ACT_GELU) {
      if (!preact) {
        set_error("linear backward: GELU gradient needs the pre-activation buffer");
        return FFK_ERR_UNSUPPORTED;
      }
      ref = preact;
    }
    FFK_REQUIRE(ref, "linear backward: activation gradient needs `output`");
    FFK_REQUIRE(aligned16(output_grad) && aligned16(ref), "linear backward: unaligned tensor");
    FFK_DISPATCH_FLOAT(dtype, T, {
      unsigned grid = stream_grid(n_out / Vec16<T>::N + 1, 256, 8);
      act_backward_kernel<T><<<grid, 256, 0, s>>>(reinterpret_cast<T *>(output_grad),
                                                  reinterpret_cast<T const *>(ref), n_out,
                                                  activation);
    });
    FFK_LAUNCH_CHECK();
  }

  {  // weight gradient: dW[o][i] += sum_b dY[b][o] X[b][i]
    gemm::Problem p{output_grad, input, out_dim, in_dim, 0, 0, out_dim, in_dim, batch_size, 1,
                    dtype, dtype, true, true};
    gemm::Epilogue e{weight_grad, in_dim, 0, nullptr, gemm::ACT_NONE, 1, 1.f,
                     l2_lambda, l2_lambda != 0.f ? weight : nullptr};
    int rc = run_gemm(p, e, s, path);
    if (rc) return rc;
  }
  if (bias_grad) {
    FFK_DISPATCH_FLOAT(dtype, T, {
      constexpr int VN = Vec16<T>::N;
      int splits = workspace ? int(std::min<size_t>(bias_grad_max_splits(out_dim),
                                                    workspace_bytes / (size_t(out_dim) * 4)))
                             : 0;
      splits = std::min(splits, (batch_size + 63) / 64);
      if (splits >= 2 && out_dim % VN == 0 && aligned16(output_grad)) {
        int rps = (batch_size + splits - 1) / splits;
        dim3 grid((out_dim / VN + 31) / 32, splits);
        column_sum_partial_kernel<T><<<grid, 256, 0, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<float *>(workspace),
            batch_size, out_dim, rps);
        column_sum_finish_kernel<T><<<(out_dim + 255) / 256, 256, 0, s>>>(
            reinterpret_cast<float const *>(workspace), reinterpret_cast<T *>(bias_grad), out_dim,
            splits);
      } else {
        column_sum_accumulate_kernel<T><<<(out_dim + 31) / 32, 256, 0, s>>>(
            reinterpret_cast<T const *>(output_grad), reinterpret_cast<T *>(bias_grad),
            batch_size, out_dim);
      }
    });
    FFK_LAUNCH_CHECK();
  }
  if (input_grad) {  // data gradient: dX[b][i] (+)= sum_o dY[b][o] W[o][i]
    gemm::Problem p{output_grad, weight, out_dim, in_dim, 0, 0, batch_size, in_dim, out_dim, 1,
                    dtype, dtype, false, true};
    gemm::Epilogue e{input_grad, in_dim, 0, nullptr, gemm::ACT_NONE, fu.accumulate_input_grad ? 1 : 0,
                     1.f, 0.f, nullptr};
    if (fu.input_gate != nullptr) {
      gemm::Epilogue g = e;
      g.r = fu.input_gate;
      g.r_mode = 1;
      g.lambda = fu.input_gate_scale;
      if ((path == FFK_GEMM_AUTO || path == FFK_GEMM_TENSOR) && gemm::fused_eligible(p, g)) {
        int rc = gemm::launch_tc(p, g, s);
        if (rc) return rc;
      } else {
        if (fu.accumulate_input_grad) {
          set_error("linear backward: a gated data gradient that accumulates needs the tensor-core "
                    "epilogue (16-bit tensors, in_dim a multiple of 32, 16-byte aligned rows)");
          return FFK_ERR_UNSUPPORTED;
        }
        int rc = run_gemm(p, e, s, path);
        if (rc) return rc;
        size_t n_in = size_t(in_dim) * batch_size;
        FFK_DISPATCH_FLOAT(dtype, T, {
          gate_kernel<T><<<stream_grid(n_in / Vec16<T>::N + 1, 256, 8), 256, 0, s>>>(
              reinterpret_cast<T *>(input_grad), reinterpret_cast<T const *>(fu.input_gate), n_in,
              fu.input_gate_scale);
        });
        FFK_LAUNCH_CHECK();
      }
    } else {
      int rc = run_gemm(p, e, s, path);
      if (rc) return rc;
    }
  }
  return FFK_OK;
}
}  // namespace

extern "C" {

int ffk_linear_backward(void *stream, void const *input, void *input_grad, void const *output,
                        void *output_grad, void const *weight, void *weight_grad,
                        void *bias_grad, int in_dim, int out_dim, int batch_size, int dtype,
                        int activation, float l2_lambda, void const *preact, int path,
                        void *workspace, size_t workspace_bytes) {
  ffk_linear_backward_fusion_t plain{};
  plain.accumulate_input_grad = 1;
  return linear_backward_impl(stream, input, input_grad, output, output_grad, weight, weight_grad, bias_grad,
                              in_dim, out_dim, batch_size, dtype, activation, l2_lambda, preact, path,
                              workspace, workspace_bytes, plain);
}

int ffk_linear_backward_fused(void *stream, void const *input, void *input_grad, void const *output,
                              void *output_grad, void const *weight, void *weight_grad,
                              void *bias_grad, int in_dim, int out_dim, int batch_size, int dtype,
                              int activation, float l2_lambda, void const *preact, int path,
                              void *workspace, size_t workspace_bytes,
                              ffk_linear_backward_fusion_t const *fusion_host) {
  FFK_REQUIRE(fusion_host, "linear backward (fused): null fusion record");
  return linear_backward_impl(stream, input, input_grad, output, output_grad, weight, weight_grad, bias_grad,
                              in_dim, out_dim, batch_size, dtype, activation, l2_lambda, preact, path,
                              workspace, workspace_bytes, *fusion_host);
}

int ffk_linear_forward_fused(void *stream, void const *input, void *output, void const *weight,
                             void const *bias, int in_dim, int out_dim, int batch_size, int dtype,
                             int activation, int path, ffk_linear_forward_fusion_t const *fusion_host) {
  FFK_REQUIRE(fusion_host, "linear forward (fused): null fusion record");
  ffk_linear_forward_fusion_t const &fu = *fusion_host;
  FFK_REQUIRE(in_dim > 0 && out_dim > 0 && batch_size >= 0, "linear forward: bad extents");
  if (batch_size == 0) return FFK_OK;
  FFK_REQUIRE(input && output && weight, "linear forward: null tensor");
  FFK_REQUIRE(fu.dropout_rate >= 0.f && fu.dropout_rate <= 1.f, "linear forward: dropout rate %f outside [0,1]",
              fu.dropout_rate);
  bool const drop = fu.dropout_rate > 0.f;
  cudaStream_t s = as_stream(stream);
  gemm::Problem p{input, weight, in_dim, in_dim, 0, 0, batch_size, out_dim, in_dim, 1,
                  dtype, dtype, false, false};
  gemm::Epilogue e{output, out_dim, 0, bias, activation, 0, 1.f, 0.f, nullptr};
  gemm::Epilogue f = e;
  f.aux_out = fu.preact_out;
  if (drop) {
    f.drop_on = 1;
    f.drop_mask = reinterpret_cast<unsigned *>(fu.dropout_mask);
    f.drop_thresh = dropout_thresh16(fu.dropout_rate);
    f.drop_scale = fu.dropout_rate < 1.f ? 1.f / (1.f - fu.dropout_rate) : 0.f;
    f.drop_seed = fu.dropout_seed;
    f.drop_offset = fu.dropout_offset;
  }
  if (fu.residual) {
    f.r = fu.residual;
    f.lambda = 1.f;
    f.r_mode = 0;
  }
  if ((path == FFK_GEMM_AUTO || path == FFK_GEMM_TENSOR) && gemm::fused_eligible(p, f))
    return gemm::launch_tc(p, f, s);   // one kernel: product, bias, activation, dropout, residual
  // the same result from the unfused chain (any dtype, any shape)
  size_t n_out = size_t(out_dim) * batch_size;
  if (fu.preact_out) {
    gemm::Epilogue pre = e;
    pre.d = fu.preact_out;
    pre.act = gemm::ACT_NONE;
    int rc = run_gemm(p, pre, s, path);
    if (rc) return rc;
  }
  int rc = run_gemm(p, e, s, path);
  if (rc) return rc;
  if (drop) {
    FFK_REQUIRE(fu.dropout_mask, "linear forward (fused): the unfused chain needs a dropout mask buffer");
    rc = ffk_dropout_forward(stream, output, output, fu.dropout_mask, n_out, fu.dropout_rate, fu.dropout_seed,
                             fu.dropout_offset, dtype);
    if (rc) return rc;
  }
  if (fu.residual) rc = ffk_accumulate(stream, output, fu.residual, n_out, dtype);
  return rc;
}

size_t ffk_linear_backward_workspace_bytes(int out_dim) {
  return out_dim > 0 ? size_t(bias_grad_max_splits(out_dim)) * size_t(out_dim) * sizeof(float) : 0;
}

int ffk_batch_matmul_forward(void *stream, void *output, void const *a_input, void const *b_input,
                             int m, int n, int k, int batch, int a_seq_length_dim,
                             int b_seq_length_dim, int seq_length, int dtype, int path) {
  FFK_REQUIRE(m >= 0 && n >= 0 && k >= 0 && batch >= 0, "batch_matmul: negative extent");
  long long const lda = k, ldb = m, ldo = m;
  long long const sa = 1LL * n * k, sb = 1LL * k * m, so = 1LL * n * m;
  // sequence-length truncation, following the reference's rules
  if (a_seq_length_dim == 0 && seq_length >= 0) {
    FFK_REQUIRE(seq_length <= k && b_seq_length_dim == 1, "batch_matmul: bad seq_length for k");
    k = seq_length;
  } else if (a_seq_length_dim == 1 && seq_length >= 0) {
    FFK_REQUIRE(seq_length <= n, "batch_matmul: seq_length > n");
    n = seq_length;
  } else {
    FFK_REQUIRE(a_seq_length_dim < 0 || seq_length < 0, "batch_matmul: a_seq_length_dim must be 0/1");
  }
  if (b_seq_length_dim == 0 && seq_length >= 0) {
    FFK_REQUIRE(seq_length <= m, "batch_matmul: seq_length > m");
    m = seq_length;
  } else if (b_seq_length_dim == 1 && seq_length >= 0) {
    FFK_REQUIRE(a_seq_length_dim == 0 && k == seq_length, "batch_matmul: inconsistent seq dims");
  } else {
    FFK_REQUIRE(b_seq_length_dim < 0 || seq_length < 0, "batch_matmul: b_seq_length_dim must be 0/1");
  }
  if (m == 0 || n == 0 || batch == 0) return FFK_OK;
  FFK_REQUIRE(output && a_input && b_input, "batch_matmul: null tensor");
  gemm::Problem p{a_input, b_input, lda, ldb, sa, sb, n, m, k, batch, dtype, dtype, false, true};
  gemm::Epilogue e{output, ldo, so, nullptr, gemm::ACT_NONE, 0, 1.f, 0.f, nullptr};
  return run_gemm(p, e, as_stream(stream), path);
}

int ffk_batch_matmul_backward(void *stream, void const *o, void const *o_grad, void const *a,
                              void *a_grad, void const *b, void *b_grad, int m, int n, int k,
                              int batch, int dtype, int path) {
  (void)o;
  FFK_REQUIRE(m >= 0 && n >= 0 && k >= 0 && batch >= 0, "batch_matmul: negative extent");
  if (m == 0 || n == 0 || k == 0 || batch == 0) return FFK_OK;
  FFK_REQUIRE(o_grad && a && a_grad && b && b_grad, "batch_matmul backward: null tensor");
  cudaStream_t s = as_stream(stream);
  long long const sa = 1LL * n * k, sb = 1LL * k * m, so = 1LL * n * m;
  {  // dA (n x k) += dO (n x m) * B^T   [B stored k x m: k-major over m for N = k]
    gemm::Problem p{o_grad, b, m, m, so, sb, n, k, m, batch, dtype, dtype, false, false};
    gemm::Epilogue e{a_grad, k, sa, nullptr, gemm::ACT_NONE, 1, 1.f, 0.f, nullptr};
    int rc = run_gemm(p, e, s, path);
    if (rc) return rc;
  }
  {  // dB (k x m) += A^T (k x n) * dO (n x m)
    gemm::Problem p{a, o_grad, k, m, sa, so, k, m, n, batch, dtype, dtype, true, true};
    gemm::Epilogue e{b_grad, m, sb, nullptr, gemm::ACT_NONE, 1, 1.f, 0.f, nullptr};
    int rc = run_gemm(p, e, s, path);
    if (rc) return rc;
  }
  return FFK_OK;
}

}  // extern "C"
