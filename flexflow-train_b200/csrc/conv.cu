// Conv2D as im2col + the GEMM engine (implicit-GEMM lowering, NCHW fp32): one column panel
// [K = in_c/g * kh * kw][out_h * out_w] per sample and group, products
//   forward      Y[g]  (out_c/g x HW)  = W[g] (out_c/g x K)  * col (K x HW)
//   filter grad  dW[g] (out_c/g x K)  += dY[g] (out_c/g x HW) * col^T
//   data grad    dcol  (K x HW)        = W[g]^T * dY[g],  then col2im accumulates into dX.
#include "gemm_sm100.cuh"

namespace ffk {
namespace {

constexpr int kThreads = 256;

struct CG {
  int n, ci, ih, iw, co, oh, ow, kh, kw, ph, pw, sh, sw, g;
  int cig, cog;       // channels per group
  long long K, HW;    // panel extents
};

int make_cg(ffk_conv2d_config_t const *c, CG &g) {
  FFK_REQUIRE(c, "conv2d: null config");
  FFK_REQUIRE(c->n >= 0 && c->in_c > 0 && c->in_h > 0 && c->in_w > 0 && c->out_c > 0 && c->kernel_h > 0 &&
                  c->kernel_w > 0 && c->stride_h > 0 && c->stride_w > 0 && c->pad_h >= 0 && c->pad_w >= 0 &&
                  c->groups > 0,
              "conv2d: bad geometry");
  FFK_REQUIRE(c->in_c % c->groups == 0 && c->out_c % c->groups == 0, "conv2d: channels must divide by groups");
  int oh = (c->in_h + 2 * c->pad_h - c->kernel_h) / c->stride_h + 1;
  int ow = (c->in_w + 2 * c->pad_w - c->kernel_w) / c->stride_w + 1;
  FFK_REQUIRE(oh == c->out_h && ow == c->out_w, "conv2d: output %dx%d does not match geometry (%dx%d expected)",
              c->out_h, c->out_w, oh, ow);
  g = CG{c->n, c->in_c, c->in_h, c->in_w, c->out_c, oh, ow, c->kernel_h, c->kernel_w, c->pad_h, c->pad_w,
         c->stride_h, c->stride_w, c->groups, c->in_c / c->groups, c->out_c / c->groups, 0, 0};
  g.K = (long long)g.cig * g.kh * g.kw;
  g.HW = (long long)oh * ow;
  return FFK_OK;
}

// col[(c*kh + a)*kw + b][ho*ow + wo] = x[group channel c][ho*sh - ph + a][wo*sw - pw + b] (0 outside)
__global__ void __launch_bounds__(kThreads)
    im2col_kernel(float const *__restrict__ x, float *__restrict__ col, CG g, int group) {
  long long const total = g.K * g.HW, gs = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gs) {
    long long k = i / g.HW, p = i - k * g.HW;
    int b = int(k % g.kw), a = int((k / g.kw) % g.kh), c = int(k / ((long long)g.kw * g.kh));
    int ho = int(p / g.ow), wo = int(p % g.ow);
    int h = ho * g.sh - g.ph + a, w = wo * g.sw - g.pw + b;
    float v = 0.f;
    if (h >= 0 && h < g.ih && w >= 0 && w < g.iw)
      v = x[((long long)(group * g.cig + c) * g.ih + h) * g.iw + w];
    col[i] = v;
  }
}

// dX[c][h][w] += sum over the kernel taps that read this pixel of dcol (gather form: no atomics)
__global__ void __launch_bounds__(kThreads)
    col2im_accumulate_kernel(float const *__restrict__ dcol, float *__restrict__ dx, CG g, int group) {
  long long const total = (long long)g.cig * g.ih * g.iw, gs = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gs) {
    int w = int(i % g.iw), h = int((i / g.iw) % g.ih), c = int(i / ((long long)g.iw * g.ih));
    float acc = 0.f;
    for (int a = 0; a < g.kh; ++a) {
      int hn = h + g.ph - a;
      if (hn < 0 || hn % g.sh) continue;
      int ho = hn / g.sh;
      if (ho >= g.oh) continue;
      for (int b = 0; b < g.kw; ++b) {
        int wn = w + g.pw - b;
        if (wn < 0 || wn % g.sw) continue;
        int wo = wn / g.sw;
        if (wo >= g.ow) continue;
        acc += dcol[(((long long)c * g.kh + a) * g.kw + b) * g.HW + (long long)ho * g.ow + wo];
      }
    }
    dx[((long long)(group * g.cig + c) * g.ih + h) * g.iw + w] += acc;
  }
}

__global__ void __launch_bounds__(kThreads)
    bias_relu_kernel(float *__restrict__ y, float const *__restrict__ bias, long long total, long long hw,
                     int co, bool relu) {
  long long const gs = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gs) {
    float v = y[i];
    if (bias) v += bias[(i / hw) % co];
    if (relu) v = fmaxf(v, 0.f);
    y[i] = v;
  }
}
__global__ void __launch_bounds__(kThreads)
    relu_grad_kernel(float *__restrict__ dy, float const *__restrict__ y, long long total) {
  long long const gs = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gs)
    dy[i] = y[i] > 0.f ? dy[i] : 0.f;
}
// db[c] += sum over n, h, w of dY: one CTA per channel
__global__ void __launch_bounds__(kThreads)
    bias_grad_kernel(float const *__restrict__ dy, float *__restrict__ db, int n, int co, long long hw) {
  __shared__ float red[32];
  int const c = blockIdx.x;
  float acc = 0.f;
  for (long long i = threadIdx.x; i < (long long)n * hw; i += blockDim.x) {
    long long b = i / hw, p = i - b * hw;
    acc += dy[(b * co + c) * hw + p];
  }
  float t = block_sum(acc, red);
  if (threadIdx.x == 0) db[c] += t;
}

int mm(cudaStream_t s, void const *a, long long lda, bool a_mn, void const *b, long long ldb, bool b_mn,
       void *d, long long ldd, long long M, long long N, long long K, bool accumulate) {
  gemm::Problem p{a, b, lda, ldb, 0, 0, (int)M, (int)N, (int)K, 1, FFK_FLOAT, FFK_FLOAT, a_mn, b_mn};
  gemm::Epilogue e{d, ldd, 0, nullptr, gemm::ACT_NONE, accumulate ? 1 : 0, 1.f, 0.f, nullptr};
  return gemm::launch(p, e, s);
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

size_t ffk_conv2d_workspace_bytes(ffk_conv2d_config_t const *cfg_host) {
  CG g;
  if (make_cg(cfg_host, g)) return 0;
  return size_t(g.K) * size_t(g.HW) * sizeof(float);
}

int ffk_conv2d_forward(void *stream, ffk_conv2d_config_t const *cfg_host, float const *input,
                       float *output, float const *filter, float const *bias, int relu,
                       void *workspace, size_t workspace_bytes) {
  CG g;
  int rc = make_cg(cfg_host, g);
  if (rc) return rc;
  if (g.n == 0) return FFK_OK;
  FFK_REQUIRE(input && output && filter && workspace, "conv2d forward: null tensor");
  size_t panel = size_t(g.K) * size_t(g.HW) * sizeof(float);
  FFK_REQUIRE(workspace_bytes >= panel, "conv2d forward: workspace of %zu bytes < %zu needed for one im2col panel",
              workspace_bytes, panel);
  cudaStream_t s = as_stream(stream);
  float *col = static_cast<float *>(workspace);
  for (int b = 0; b < g.n; ++b) {
    float const *xb = input + (long long)b * g.ci * g.ih * g.iw;
    float *yb = output + (long long)b * g.co * g.HW;
    for (int grp = 0; grp < g.g; ++grp) {
      im2col_kernel<<<stream_grid(size_t(g.K * g.HW), kThreads, 8), kThreads, 0, s>>>(xb, col, g, grp);
      FFK_LAUNCH_CHECK();
      float const *wg = filter + (long long)grp * g.cog * g.K;
      if ((rc = mm(s, wg, g.K, false, col, g.HW, true, yb + (long long)grp * g.cog * g.HW, g.HW, g.cog, g.HW, g.K, false)))
        return rc;
    }
  }
  if (bias || relu) {
    long long total = (long long)g.n * g.co * g.HW;
    bias_relu_kernel<<<stream_grid(size_t(total), kThreads, 8), kThreads, 0, s>>>(output, bias, total, g.HW, g.co, relu != 0);
    FFK_LAUNCH_CHECK();
  }
  return FFK_OK;
}

int ffk_conv2d_backward(void *stream, ffk_conv2d_config_t const *cfg_host, float const *input,
                        float *input_grad, float const *output, float *output_grad,
                        float const *filter, float *filter_grad, float *bias_grad, int relu,
                        void *workspace, size_t workspace_bytes) {
  CG g;
  int rc = make_cg(cfg_host, g);
  if (rc) return rc;
  if (g.n == 0) return FFK_OK;
  FFK_REQUIRE(input && output_grad && filter && filter_grad && workspace, "conv2d backward: null tensor");
  FFK_REQUIRE(!relu || output, "conv2d backward: relu needs `output`");
  size_t panel = size_t(g.K) * size_t(g.HW) * sizeof(float);
  FFK_REQUIRE(workspace_bytes >= panel, "conv2d backward: workspace of %zu bytes < %zu needed", workspace_bytes, panel);
  cudaStream_t s = as_stream(stream);
  float *col = static_cast<float *>(workspace);
  long long total = (long long)g.n * g.co * g.HW;
  if (relu) {
    relu_grad_kernel<<<stream_grid(size_t(total), kThreads, 8), kThreads, 0, s>>>(output_grad, output, total);
    FFK_LAUNCH_CHECK();
  }
  for (int b = 0; b < g.n; ++b) {
    float const *xb = input + (long long)b * g.ci * g.ih * g.iw;
    float const *dyb = output_grad + (long long)b * g.co * g.HW;
    for (int grp = 0; grp < g.g; ++grp) {
      float const *dyg = dyb + (long long)grp * g.cog * g.HW;
      // filter gradient: dW[g] += dY[g] col^T
      im2col_kernel<<<stream_grid(size_t(g.K * g.HW), kThreads, 8), kThreads, 0, s>>>(xb, col, g, grp);
      FFK_LAUNCH_CHECK();
      if ((rc = mm(s, dyg, g.HW, false, col, g.HW, false, filter_grad + (long long)grp * g.cog * g.K, g.K, g.cog, g.K, g.HW, true)))
        return rc;
      if (input_grad) {
        // data gradient: dcol = W[g]^T dY[g], gathered back into dX
        float const *wg = filter + (long long)grp * g.cog * g.K;
        if ((rc = mm(s, wg, g.K, true, dyg, g.HW, true, col, g.HW, g.K, g.HW, g.cog, false))) return rc;
        float *dxb = input_grad + (long long)b * g.ci * g.ih * g.iw;
        col2im_accumulate_kernel<<<stream_grid(size_t((long long)g.cig * g.ih * g.iw), kThreads, 8), kThreads, 0, s>>>(col, dxb, g, grp);
        FFK_LAUNCH_CHECK();
      }
    }
  }
  if (bias_grad) {
    bias_grad_kernel<<<g.co, kThreads, 0, s>>>(output_grad, bias_grad, g.n, g.co, g.HW);
    FFK_LAUNCH_CHECK();
  }
  return FFK_OK;
}

}  // extern "C"
