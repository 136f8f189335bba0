// oracle/cudnn_ref/ops_golden.cu — TEST INFRASTRUCTURE (golden-vector generator), not product.
//
// For the operators whose arithmetic the reference delegates to cuDNN, this program issues the
// SAME cuDNN calls with the SAME descriptor settings the reference uses (cited per case) on
// seeded inputs on the GPU box and dumps inputs and results. tools/make_ops_golden.py packs
// them into tests/golden/cudnn_ops.npz; the CPU oracle and the sm_100a kernels are then both
// checked against what the reference's arithmetic provider actually returns.
//
// Build (here):  nvcc -O2 -o oracle/_build/ops_golden oracle/cudnn_ref/ops_golden.cu -lcudnn
// Run (GPU box): oracle/_build/ops_golden <outdir>
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

static unsigned long long rng_state = 0x9E3779B97F4A7C15ULL;
static float urand() {
  rng_state ^= rng_state >> 12;
  rng_state ^= rng_state << 25;
  rng_state ^= rng_state >> 27;
  unsigned long long r = rng_state * 2685821657736338717ULL;
  return (float)((r >> 40) / 8388608.0 - 1.0);
}
static std::vector<float> rnd(size_t n, float lo = -1.f, float hi = 1.f) {
  std::vector<float> v(n);
  for (auto &x : v) x = lo + (hi - lo) * (urand() * 0.5f + 0.5f);
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

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  g_dir = argv[1];
  g_meta = fopen((g_dir + "/meta.json").c_str(), "w");
  fprintf(g_meta, "{");
  cudnnHandle_t h;
  CK(cudnnCreate(&h));
  float one = 1.f, zero = 0.f;

  {  // Softmax forward — softmax_kernels.cu:45-61: ACCURATE, MODE_CHANNEL over (n,c,h,w)
    int n = 3, c = 37, hh = 2, w = 5;
    auto x = rnd((size_t)n * c * hh * w, -3, 3);
    Dev dx(x), dy(x.size());
    auto d = nchw(n, c, hh, w);
    CK(cudnnSoftmaxForward(h, CUDNN_SOFTMAX_ACCURATE, CUDNN_SOFTMAX_MODE_CHANNEL, &one, d, dx.p, &zero, d, dy.p));
    dump("softmax_x", x, {n, c, hh, w});
    dump("softmax_y", dy.host(), {n, c, hh, w});
  }
  {  // ElementUnary RELU / SIGMOID / TANH / ELU — element_unary_kernels.cu:60-82 (coef 0.0), :259 fwd, :301 bwd (beta = 1)
    int n = 4, c = 3, hh = 5, w = 7;
    size_t sz = (size_t)n * c * hh * w;
    auto x = rnd(sz, -2, 2), gy = rnd(sz), gx0 = rnd(sz);
    dump("act_x", x, {n, c, hh, w});
    dump("act_dy", gy, {n, c, hh, w});
    dump("act_dx0", gx0, {n, c, hh, w});
    auto d = nchw(n, c, hh, w);
    cudnnActivationMode_t modes[4] = {CUDNN_ACTIVATION_RELU, CUDNN_ACTIVATION_SIGMOID, CUDNN_ACTIVATION_TANH, CUDNN_ACTIVATION_ELU};
    char const *names[4] = {"relu", "sigmoid", "tanh", "elu"};
    for (int i = 0; i < 4; ++i) {
      cudnnActivationDescriptor_t a;
      CK(cudnnCreateActivationDescriptor(&a));
      CK(cudnnSetActivationDescriptor(a, modes[i], CUDNN_PROPAGATE_NAN, 0.0));
      Dev dx(x), dy(sz), dgy(gy), dgx(gx0);
      CK(cudnnActivationForward(h, a, &one, d, dx.p, &zero, d, dy.p));
      CK(cudnnActivationBackward(h, a, &one, d, dy.p, d, dgy.p, d, dx.p, &one, d, dgx.p));
      dump(std::string("act_") + names[i] + "_y", dy.host(), {n, c, hh, w});
      dump(std::string("act_") + names[i] + "_dx", dgx.host(), {n, c, hh, w});
    }
  }
  {  // ElementBinary — element_binary_kernels.cu:141-243 (cudnnOpTensor, RHS broadcast per channel)
    int n = 4, c = 6, hh = 3, w = 5;
    size_t sz = (size_t)n * c * hh * w;
    auto a = rnd(sz), b = rnd(c);
    dump("bin_lhs", a, {n, c, hh, w});
    dump("bin_rhs", b, {1, c, 1, 1});
    auto da_ = nchw(n, c, hh, w), db_ = nchw(1, c, 1, 1);
    cudnnOpTensorOp_t ops[4] = {CUDNN_OP_TENSOR_ADD, CUDNN_OP_TENSOR_MUL, CUDNN_OP_TENSOR_MAX, CUDNN_OP_TENSOR_MIN};
    char const *names[4] = {"add", "mul", "max", "min"};
    for (int i = 0; i < 4; ++i) {
      cudnnOpTensorDescriptor_t od;
      CK(cudnnCreateOpTensorDescriptor(&od));
      CK(cudnnSetOpTensorDescriptor(od, ops[i], CUDNN_DATA_FLOAT, CUDNN_PROPAGATE_NAN));
      Dev dA(a), dB(b), dO(sz);
      CK(cudnnOpTensor(h, od, &one, da_, dA.p, &one, db_, dB.p, &zero, da_, dO.p));
      dump(std::string("bin_") + names[i], dO.host(), {n, c, hh, w});
    }
    // EW_SUB = ADD with alpha2 = -1 (:152-154)
    cudnnOpTensorDescriptor_t od;
    CK(cudnnCreateOpTensorDescriptor(&od));
    CK(cudnnSetOpTensorDescriptor(od, CUDNN_OP_TENSOR_ADD, CUDNN_DATA_FLOAT, CUDNN_PROPAGATE_NAN));
    float neg = -1.f;
    Dev dA(a), dB(b), dO(sz);
    CK(cudnnOpTensor(h, od, &one, da_, dA.p, &neg, db_, dB.p, &zero, da_, dO.p));
    dump("bin_sub", dO.host(), {n, c, hh, w});
    // backward of ADD for the broadcast side: cudnnReduceTensor ADD with beta = 1 (:263-300)
    auto gy = rnd(sz), g0 = rnd(c);
    dump("bin_dy", gy, {n, c, hh, w});
    dump("bin_drhs0", g0, {1, c, 1, 1});
    cudnnReduceTensorDescriptor_t rd;
    CK(cudnnCreateReduceTensorDescriptor(&rd));
    CK(cudnnSetReduceTensorDescriptor(rd, CUDNN_REDUCE_TENSOR_ADD, CUDNN_DATA_FLOAT, CUDNN_PROPAGATE_NAN,
                                      CUDNN_REDUCE_TENSOR_NO_INDICES, CUDNN_32BIT_INDICES));
    size_t ws = 0;
    CK(cudnnGetReductionWorkspaceSize(h, rd, da_, db_, &ws));
    void *wsp;
    CK(cudaMalloc(&wsp, ws ? ws : 16));
    Dev dGy(gy), dG(g0);
    CK(cudnnReduceTensor(h, rd, nullptr, 0, wsp, ws, &one, da_, dGy.p, &one, db_, dG.p));
    dump("bin_add_drhs", dG.host(), {1, c, 1, 1});
  }
  {  // Reduce — reduce_kernels.cu:47-65 with the intended SUM (ADD) and MEAN (AVG) reductions
    int n = 5, c = 4, hh = 6, w = 3;
    size_t sz = (size_t)n * c * hh * w;
    auto x = rnd(sz);
    dump("reduce_x", x, {n, c, hh, w});
    auto di = nchw(n, c, hh, w), dо = nchw(n, 1, 1, w);
    cudnnReduceTensorOp_t ops[2] = {CUDNN_REDUCE_TENSOR_ADD, CUDNN_REDUCE_TENSOR_AVG};
    char const *names[2] = {"reduce_sum", "reduce_mean"};
    for (int i = 0; i < 2; ++i) {
      cudnnReduceTensorDescriptor_t rd;
      CK(cudnnCreateReduceTensorDescriptor(&rd));
      CK(cudnnSetReduceTensorDescriptor(rd, ops[i], CUDNN_DATA_FLOAT, CUDNN_PROPAGATE_NAN,
                                        CUDNN_REDUCE_TENSOR_NO_INDICES, CUDNN_32BIT_INDICES));
      size_t ws = 0;
      CK(cudnnGetReductionWorkspaceSize(h, rd, di, dо, &ws));
      void *wsp;
      CK(cudaMalloc(&wsp, ws ? ws : 16));
      Dev dX(x), dY((size_t)n * w);
      CK(cudnnReduceTensor(h, rd, nullptr, 0, wsp, ws, &one, di, dX.p, &zero, dо, dY.p));
      dump(names[i], dY.host(), {n, 1, 1, w});
    }
  }
  {  // BatchNorm — batch_norm_kernels.cu:26-90: SPATIAL_PERSISTENT, factor 1.0, eps = CUDNN_BN_MIN_EPSILON,
     // backward with all alpha/beta = 1
    int n = 6, c = 5, hh = 4, w = 3;
    size_t sz = (size_t)n * c * hh * w;
    auto x = rnd(sz), sc = rnd(c, 0.5f, 1.5f), bi = rnd(c), gy = rnd(sz), gx0 = rnd(sz), gs0 = rnd(c), gb0 = rnd(c);
    auto d = nchw(n, c, hh, w), dc = nchw(1, c, 1, 1);
    Dev dX(x), dY(sz), dS(sc), dB(bi), rm(c), rv(c), sm(c), sv(c), dGy(gy), dGx(gx0), dGs(gs0), dGb(gb0);
    double eps = CUDNN_BN_MIN_EPSILON;
    CK(cudnnBatchNormalizationForwardTraining(h, CUDNN_BATCHNORM_SPATIAL_PERSISTENT, &one, &zero, d, dX.p, d, dY.p,
                                              dc, dS.p, dB.p, 1.0, rm.p, rv.p, eps, sm.p, sv.p));
    CK(cudnnBatchNormalizationBackward(h, CUDNN_BATCHNORM_SPATIAL_PERSISTENT, &one, &one, &one, &one, d, dX.p, d,
                                       dGy.p, d, dGx.p, dc, dS.p, dGs.p, dGb.p, eps, sm.p, sv.p));
    dump("bn_x", x, {n, c, hh, w}); dump("bn_scale", sc, {c}); dump("bn_bias", bi, {c});
    dump("bn_dy", gy, {n, c, hh, w}); dump("bn_dx0", gx0, {n, c, hh, w}); dump("bn_dscale0", gs0, {c}); dump("bn_dbias0", gb0, {c});
    dump("bn_y", dY.host(), {n, c, hh, w}); dump("bn_running_mean", rm.host(), {c}); dump("bn_running_var", rv.host(), {c});
    dump("bn_save_mean", sm.host(), {c}); dump("bn_save_invvar", sv.host(), {c});
    dump("bn_dx", dGx.host(), {n, c, hh, w}); dump("bn_dscale", dGs.host(), {c}); dump("bn_dbias", dGb.host(), {c});
    dump("bn_eps", {(float)eps}, {1});
  }
  {  // Pool2D — pool_2d_kernels.cu:61-135: MAX and AVERAGE_COUNT_EXCLUDE_PADDING, backward beta = 1
    int n = 2, c = 3, ih = 9, iw = 11, k = 3, s = 2, p = 1;
    int oh = (ih + 2 * p - k) / s + 1, ow = (iw + 2 * p - k) / s + 1;
    size_t isz = (size_t)n * c * ih * iw, osz = (size_t)n * c * oh * ow;
    auto x = rnd(isz), gy = rnd(osz), gx0 = rnd(isz);
    dump("pool_x", x, {n, c, ih, iw}); dump("pool_dy", gy, {n, c, oh, ow}); dump("pool_dx0", gx0, {n, c, ih, iw});
    auto di = nchw(n, c, ih, iw), dо = nchw(n, c, oh, ow);
    cudnnPoolingMode_t modes[2] = {CUDNN_POOLING_MAX, CUDNN_POOLING_AVERAGE_COUNT_EXCLUDE_PADDING};
    char const *names[2] = {"pool_max", "pool_avg"};
    for (int i = 0; i < 2; ++i) {
      cudnnPoolingDescriptor_t pd;
      CK(cudnnCreatePoolingDescriptor(&pd));
      CK(cudnnSetPooling2dDescriptor(pd, modes[i], CUDNN_PROPAGATE_NAN, k, k, p, p, s, s));
      Dev dX(x), dY(osz), dGy(gy), dGx(gx0);
      CK(cudnnPoolingForward(h, pd, &one, di, dX.p, &zero, dо, dY.p));
      CK(cudnnPoolingBackward(h, pd, &one, dо, dY.p, dо, dGy.p, di, dX.p, &one, di, dGx.p));
      dump(std::string(names[i]) + "_y", dY.host(), {n, c, oh, ow});
      dump(std::string(names[i]) + "_dx", dGx.host(), {n, c, ih, iw});
    }
  }
  {  // Dropout — dropout_kernels.cu:24-94: the mask is cuDNN's own Philox stream; what is pinned is
     // the value rule y in {0, x/(1-p)} and that backward reuses the forward mask
    int n = 1, c = 1, hh = 100, w = 100;
    size_t sz = (size_t)hh * w;
    float rate = 0.1f;
    auto x = rnd(sz, 0.5f, 1.5f), gy = rnd(sz, 0.5f, 1.5f);
    auto d = nchw(n, c, hh, w);
    cudnnDropoutDescriptor_t dd;
    CK(cudnnCreateDropoutDescriptor(&dd));
    size_t ssz = 0, rsz = 0;
    CK(cudnnDropoutGetStatesSize(h, &ssz));
    CK(cudnnDropoutGetReserveSpaceSize(d, &rsz));
    void *states, *reserve;
    CK(cudaMalloc(&states, ssz));
    CK(cudaMalloc(&reserve, rsz ? rsz : 16));
    CK(cudnnSetDropoutDescriptor(dd, h, rate, states, ssz, 12345ull));
    Dev dX(x), dY(sz), dGy(gy), dGx(sz);
    CK(cudnnDropoutForward(h, dd, d, dX.p, d, dY.p, reserve, rsz));
    CK(cudnnDropoutBackward(h, dd, d, dGy.p, d, dGx.p, reserve, rsz));
    dump("dropout_x", x, {hh, w}); dump("dropout_y", dY.host(), {hh, w});
    dump("dropout_dy", gy, {hh, w}); dump("dropout_dx", dGx.host(), {hh, w});
    dump("dropout_rate", {rate}, {1});
  }
  CK(cudaDeviceSynchronize());
  fprintf(g_meta, "}\n");
  fclose(g_meta);
  printf("ok cudnn %zu\n", cudnnGetVersion());
  return 0;
}
