// oracle/cudnn_ref/conv_golden.cu — TEST INFRASTRUCTURE (golden-vector generator), not product.
//
// Conv2D as the reference runs it (lib/kernels/src/cuda/ops/conv_2d_kernels.cu): NCHW fp32 tensors,
// cudnnSetConvolution2dDescriptor(pad, stride, dilation 1, CUDNN_CROSS_CORRELATION) (:178-186),
// cudnnSetConvolutionGroupCount (:187-189); forward = cudnnConvolutionForward (alpha 1, beta 0) +
// cudnnAddTensor(bias) + cudnnActivationForward(RELU) (:276-312); backward = reluBackward on dY in
// place (:329-344), cudnnConvolutionBackwardFilter / BackwardBias / BackwardData all with
// alpha = beta = 1, i.e. ACCUMULATING onto the gradients (:345-388).
//
// Two settings differ from the reference on purpose, both so that the result is a function of the
// inputs alone at fp32 accuracy: the math type is CUDNN_FMA_MATH (the reference allows tensor-op
// down-conversion of fp32, :191-197, which is good to ~1e-3 only), and the algorithms are fixed
// deterministic ones instead of the fastest found by cudnnFind*AlgorithmEx (:8-114).
//
// Build (here):  nvcc -O2 -o oracle/_build/conv_golden oracle/cudnn_ref/conv_golden.cu -lcudnn
// Run (GPU box): oracle/_build/conv_golden <outdir>   then tools/make_ops_golden.py <outdir> cudnn_conv.npz
#include <cuda_runtime.h>
#include <cudnn.h>

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#define CK(x)                                                                            \
  do {                                                                                   \
    auto _s = (x);                                                                       \
    if (_s != 0) {                                                                       \
      fprintf(stderr, "failure %d at %s:%d: %s\n", (int)_s, __FILE__, __LINE__, #x);     \
      exit(1);                                                                           \
    }                                                                                    \
  } while (0)

static unsigned long long rng_state = 0xD1B54A32D192ED03ULL;
static float urand() {
  rng_state ^= rng_state >> 12;
  rng_state ^= rng_state << 25;
  rng_state ^= rng_state >> 27;
  unsigned long long r = rng_state * 2685821657736338717ULL;
  return (float)((r >> 40) / 8388608.0 - 1.0);
}
static std::vector<float> rnd(size_t n, float scale = 1.f) {
  std::vector<float> v(n);
  for (auto &x : v) x = scale * urand();
  return v;
}

static std::string g_dir;
static FILE *g_meta;
static bool g_first = true;
static void dump(std::string const &name, std::vector<float> const &v, std::vector<int> const &shape) {
  FILE *f = fopen((g_dir + "/" + name + ".bin").c_str(), "wb");
  fwrite(v.data(), sizeof(float), v.size(), f);
  fclose(f);
  fprintf(g_meta, "%s\"%s\": [", g_first ? "" : ",\n", name.c_str());
  for (size_t i = 0; i < shape.size(); ++i) fprintf(g_meta, "%d%s", shape[i], i + 1 < shape.size() ? ", " : "");
  fprintf(g_meta, "]");
  g_first = false;
}
struct Dev {
  float *p;
  size_t n;
  explicit Dev(std::vector<float> const &h) : n(h.size()) {
    CK(cudaMalloc(&p, n * 4));
    CK(cudaMemcpy(p, h.data(), n * 4, cudaMemcpyHostToDevice));
  }
  explicit Dev(size_t n_) : n(n_) {
    CK(cudaMalloc(&p, n * 4));
    CK(cudaMemset(p, 0, n * 4));
  }
  std::vector<float> host() const {
    std::vector<float> h(n);
    CK(cudaMemcpy(h.data(), p, n * 4, cudaMemcpyDeviceToHost));
    return h;
  }
};
static cudnnTensorDescriptor_t nchw(int n, int c, int h, int w) {
  cudnnTensorDescriptor_t d;
  CK(cudnnCreateTensorDescriptor(&d));
  CK(cudnnSetTensor4dDescriptor(d, CUDNN_TENSOR_NCHW, CUDNN_DATA_FLOAT, n, c, h, w));
  return d;
}

__global__ void relu_backward(float *dy, float const *y, int n) {  // reluBackward, cuda_helper.cu:142-148
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dy[i] = y[i] > 0.f ? dy[i] : 0.f;
}

struct Case {
  char const *name;
  int n, ci, ih, iw, co, k, s, p, g, relu;
};

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  g_dir = argv[1];
  g_meta = fopen((g_dir + "/meta.json").c_str(), "w");
  fprintf(g_meta, "{");
  cudnnHandle_t h;
  CK(cudnnCreate(&h));
  float const one = 1.f, zero = 0.f;
  // the first two geometries are tests/test_shape_gpu.py's; the last has 1 x 1 filters and more channels
  Case cases[] = {{"conv_a", 2, 3, 9, 11, 8, 3, 1, 1, 1, 1},
                  {"conv_b", 3, 4, 8, 8, 6, 3, 2, 0, 2, 0},
                  {"conv_c", 2, 6, 7, 5, 6, 5, 2, 2, 3, 1},
                  {"conv_d", 1, 16, 12, 12, 32, 1, 1, 0, 1, 0}};
  for (Case const &c : cases) {
    int const oh = (c.ih + 2 * c.p - c.k) / c.s + 1, ow = (c.iw + 2 * c.p - c.k) / c.s + 1;
    size_t const xs = (size_t)c.n * c.ci * c.ih * c.iw, ys = (size_t)c.n * c.co * oh * ow;
    size_t const ws = (size_t)c.co * (c.ci / c.g) * c.k * c.k;
    auto x = rnd(xs), w = rnd(ws, 1.f / sqrtf((float)(c.ci / c.g * c.k * c.k))), b = rnd(c.co);
    auto dy = rnd(ys), dx0 = rnd(xs, 0.5f), dw0 = rnd(ws, 0.5f), db0 = rnd(c.co, 0.5f);
    std::string const nm = c.name;
    dump(nm + "_geo", {(float)c.n, (float)c.ci, (float)c.ih, (float)c.iw, (float)c.co, (float)oh, (float)ow, (float)c.k,
                       (float)c.k, (float)c.p, (float)c.p, (float)c.s, (float)c.s, (float)c.g, (float)c.relu}, {15});
    dump(nm + "_x", x, {c.n, c.ci, c.ih, c.iw});
    dump(nm + "_w", w, {c.co, c.ci / c.g, c.k, c.k});
    dump(nm + "_b", b, {c.co});
    dump(nm + "_dy", dy, {c.n, c.co, oh, ow});
    dump(nm + "_dx0", dx0, {c.n, c.ci, c.ih, c.iw});
    dump(nm + "_dw0", dw0, {c.co, c.ci / c.g, c.k, c.k});
    dump(nm + "_db0", db0, {c.co});

    auto xd = nchw(c.n, c.ci, c.ih, c.iw), yd = nchw(c.n, c.co, oh, ow), bd = nchw(1, c.co, 1, 1);
    cudnnFilterDescriptor_t fd;
    CK(cudnnCreateFilterDescriptor(&fd));
    CK(cudnnSetFilter4dDescriptor(fd, CUDNN_DATA_FLOAT, CUDNN_TENSOR_NCHW, c.co, c.ci / c.g, c.k, c.k));
    cudnnConvolutionDescriptor_t cd;
    CK(cudnnCreateConvolutionDescriptor(&cd));
    CK(cudnnSetConvolution2dDescriptor(cd, c.p, c.p, c.s, c.s, 1, 1, CUDNN_CROSS_CORRELATION, CUDNN_DATA_FLOAT));
    if (c.g != 1) CK(cudnnSetConvolutionGroupCount(cd, c.g));
    CK(cudnnSetConvolutionMathType(cd, CUDNN_FMA_MATH));
    int on, oc, ohh, oww;
    CK(cudnnGetConvolution2dForwardOutputDim(cd, xd, fd, &on, &oc, &ohh, &oww));
    if (on != c.n || oc != c.co || ohh != oh || oww != ow) {
      fprintf(stderr, "%s: output shape %d %d %d %d\n", c.name, on, oc, ohh, oww);
      return 1;
    }
    Dev dX(x), dW(w), dB(b), dY(ys), dGy(dy), dGx(dx0), dGw(dw0), dGb(db0);
    void *wsp = nullptr;
    size_t const wsp_bytes = 64u << 20;
    CK(cudaMalloc(&wsp, wsp_bytes));

    // forward: first algorithm of the list that accepts this geometry
    cudnnConvolutionFwdAlgo_t const fwd[] = {CUDNN_CONVOLUTION_FWD_ALGO_IMPLICIT_GEMM,
                                             CUDNN_CONVOLUTION_FWD_ALGO_IMPLICIT_PRECOMP_GEMM,
                                             CUDNN_CONVOLUTION_FWD_ALGO_GEMM};
    bool done = false;
    for (auto a : fwd) {
      size_t need = 0;
      if (cudnnGetConvolutionForwardWorkspaceSize(h, xd, fd, cd, yd, a, &need) != CUDNN_STATUS_SUCCESS || need > wsp_bytes)
        continue;
      if (cudnnConvolutionForward(h, &one, xd, dX.p, fd, dW.p, cd, a, wsp, wsp_bytes, &zero, yd, dY.p) ==
          CUDNN_STATUS_SUCCESS) {
        done = true;
        break;
      }
    }
    if (!done) { fprintf(stderr, "%s: no forward algorithm\n", c.name); return 1; }
    CK(cudnnAddTensor(h, &one, bd, dB.p, &one, yd, dY.p));
    if (c.relu) {
      cudnnActivationDescriptor_t ad;
      CK(cudnnCreateActivationDescriptor(&ad));
      CK(cudnnSetActivationDescriptor(ad, CUDNN_ACTIVATION_RELU, CUDNN_NOT_PROPAGATE_NAN, 0.0));
      CK(cudnnActivationForward(h, ad, &one, yd, dY.p, &zero, yd, dY.p));
    }
    dump(nm + "_y", dY.host(), {c.n, c.co, oh, ow});

    // backward
    if (c.relu) relu_backward<<<(unsigned)((ys + 255) / 256), 256>>>(dGy.p, dY.p, (int)ys);
    cudnnConvolutionBwdFilterAlgo_t const bwf[] = {CUDNN_CONVOLUTION_BWD_FILTER_ALGO_1, CUDNN_CONVOLUTION_BWD_FILTER_ALGO_0,
                                                   CUDNN_CONVOLUTION_BWD_FILTER_ALGO_3};
    done = false;
    for (auto a : bwf) {
      size_t need = 0;
      if (cudnnGetConvolutionBackwardFilterWorkspaceSize(h, xd, yd, cd, fd, a, &need) != CUDNN_STATUS_SUCCESS ||
          need > wsp_bytes)
        continue;
      Dev trial(dw0);
      if (cudnnConvolutionBackwardFilter(h, &one, xd, dX.p, yd, dGy.p, cd, a, wsp, wsp_bytes, &one, fd, trial.p) ==
          CUDNN_STATUS_SUCCESS) {
        CK(cudaMemcpy(dGw.p, trial.p, ws * 4, cudaMemcpyDeviceToDevice));
        done = true;
        break;
      }
    }
    if (!done) { fprintf(stderr, "%s: no filter-gradient algorithm\n", c.name); return 1; }
    CK(cudnnConvolutionBackwardBias(h, &one, yd, dGy.p, &one, bd, dGb.p));
    cudnnConvolutionBwdDataAlgo_t const bwd[] = {CUDNN_CONVOLUTION_BWD_DATA_ALGO_1, CUDNN_CONVOLUTION_BWD_DATA_ALGO_0};
    done = false;
    for (auto a : bwd) {
      size_t need = 0;
      if (cudnnGetConvolutionBackwardDataWorkspaceSize(h, fd, yd, cd, xd, a, &need) != CUDNN_STATUS_SUCCESS ||
          need > wsp_bytes)
        continue;
      Dev trial(dx0);
      if (cudnnConvolutionBackwardData(h, &one, fd, dW.p, yd, dGy.p, cd, a, wsp, wsp_bytes, &one, xd, trial.p) ==
          CUDNN_STATUS_SUCCESS) {
        CK(cudaMemcpy(dGx.p, trial.p, xs * 4, cudaMemcpyDeviceToDevice));
        done = true;
        break;
      }
    }
    if (!done) { fprintf(stderr, "%s: no data-gradient algorithm\n", c.name); return 1; }
    CK(cudaDeviceSynchronize());
    dump(nm + "_dy_after", dGy.host(), {c.n, c.co, oh, ow});
    dump(nm + "_dw", dGw.host(), {c.co, c.ci / c.g, c.k, c.k});
    dump(nm + "_db", dGb.host(), {c.co});
    dump(nm + "_dx", dGx.host(), {c.n, c.ci, c.ih, c.iw});
    CK(cudaFree(wsp));
  }
  CK(cudaDeviceSynchronize());
  fprintf(g_meta, "}\n");
  fclose(g_meta);
  printf("ok cudnn %zu\n", cudnnGetVersion());
  return 0;
}
