// oracle/cudnn_ref/mha_golden.cu — TEST INFRASTRUCTURE (golden-vector generator), not product.
//
// The reference's MultiHeadAttention kernels are three cuDNN calls
// (lib/kernels/src/cuda/ops/attention_kernels.cu:255 cudnnMultiHeadAttnForward,
//  :292 cudnnMultiHeadAttnBackwardData, :316 cudnnMultiHeadAttnBackwardWeights) on a
// descriptor built at :106-125 (ALL_TO_ONE, smScalar 1.0, fp32 data and math, no dropout,
// no biases, window [0, kvSeqLength)) with BATCH,TIME,BEAM,VECT axes (:134-137).
// cuDNN is closed source, so the arithmetic cannot be restated from the reference tree; this
// program instead RUNS those calls, configured exactly as the reference configures them, on
// seeded inputs on the GPU box and records (a) the packed weight layout reported by
// cudnnGetMultiHeadAttnWeights and (b) inputs/outputs, which tools/make_mha_golden.py turns
// into tests/golden/mha_*.npz. The CPU oracle (oracle/cpu_ref.cc: oracle_mha_*) and the
// sm_100a kernels are both checked against those vectors.
//
// Build (here):  nvcc -O2 -o oracle/_build/mha_golden oracle/cudnn_ref/mha_golden.cu -lcudnn
// Run (GPU box): oracle/_build/mha_golden <outdir> N H qSize kSize vSize qProj kProj vProj oProj qoSeq kvSeq
#include <cuda_runtime.h>
#include <cudnn.h>

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#define CK(x)                                                                                 \
  do {                                                                                        \
    auto _s = (x);                                                                            \
    if (_s != 0) {                                                                            \
      fprintf(stderr, "failure %d at %s:%d: %s\n", (int)_s, __FILE__, __LINE__, #x);          \
      exit(1);                                                                                \
    }                                                                                         \
  } while (0)

static unsigned long long rng_state = 88172645463325252ULL;
static float urand() {  // xorshift64*, uniform in [-1, 1)
  rng_state ^= rng_state >> 12;
  rng_state ^= rng_state << 25;
  rng_state ^= rng_state >> 27;
  unsigned long long r = rng_state * 2685821657736338717ULL;
  return (float)((r >> 40) / 8388608.0 - 1.0);
}

static void dump(std::string const &dir, char const *name, std::vector<float> const &v) {
  std::string p = dir + "/" + name + ".bin";
  FILE *f = fopen(p.c_str(), "wb");
  fwrite(v.data(), sizeof(float), v.size(), f);
  fclose(f);
}

int main(int argc, char **argv) {
  if (argc < 13) {
    fprintf(stderr, "usage: %s outdir N H qSize kSize vSize qProj kProj vProj oProj qoSeq kvSeq\n", argv[0]);
    return 2;
  }
  std::string dir = argv[1];
  int N = atoi(argv[2]), H = atoi(argv[3]), qSize = atoi(argv[4]), kSize = atoi(argv[5]),
      vSize = atoi(argv[6]), qProj = atoi(argv[7]), kProj = atoi(argv[8]), vProj = atoi(argv[9]),
      oProj = atoi(argv[10]), qoSeq = atoi(argv[11]), kvSeq = atoi(argv[12]);

  cudnnHandle_t h;
  CK(cudnnCreate(&h));
  cudnnAttnDescriptor_t attn;
  cudnnSeqDataDescriptor_t qD, kD, vD, oD;
  CK(cudnnCreateAttnDescriptor(&attn));
  CK(cudnnCreateSeqDataDescriptor(&qD));
  CK(cudnnCreateSeqDataDescriptor(&kD));
  CK(cudnnCreateSeqDataDescriptor(&vD));
  CK(cudnnCreateSeqDataDescriptor(&oD));
  CK(cudnnSetAttnDescriptor(attn, CUDNN_ATTN_QUERYMAP_ALL_TO_ONE, H, 1.0 /*smScalar*/,
                            CUDNN_DATA_FLOAT, CUDNN_DATA_FLOAT, CUDNN_DEFAULT_MATH, NULL, NULL, qSize,
                            kSize, vSize, qProj, kProj, vProj, oProj, qoSeq, kvSeq, N, 1));
  size_t weightSize, workSize, reserveSize;
  CK(cudnnGetMultiHeadAttnBuffers(h, attn, &weightSize, &workSize, &reserveSize));

  int dimA[CUDNN_SEQDATA_DIM_COUNT];
  cudnnSeqDataAxis_t axes[CUDNN_SEQDATA_DIM_COUNT];
  axes[3] = CUDNN_SEQDATA_VECT_DIM;
  axes[2] = CUDNN_SEQDATA_BEAM_DIM;
  axes[1] = CUDNN_SEQDATA_TIME_DIM;
  axes[0] = CUDNN_SEQDATA_BATCH_DIM;
  std::vector<int> qoLen(N, qoSeq), kvLen(N, kvSeq);
  auto set = [&](cudnnSeqDataDescriptor_t d, int t, int vect, std::vector<int> &len) {
    dimA[CUDNN_SEQDATA_BEAM_DIM] = 1;
    dimA[CUDNN_SEQDATA_BATCH_DIM] = N;
    dimA[CUDNN_SEQDATA_TIME_DIM] = t;
    dimA[CUDNN_SEQDATA_VECT_DIM] = vect;
    CK(cudnnSetSeqDataDescriptor(d, CUDNN_DATA_FLOAT, CUDNN_SEQDATA_DIM_COUNT, dimA, axes, N,
                                 len.data(), NULL));
  };
  set(qD, qoSeq, qSize, qoLen);
  set(kD, kvSeq, kSize, kvLen);
  set(vD, kvSeq, vSize, kvLen);
  set(oD, qoSeq, oProj, qoLen);

  int *devQo, *devKv;
  CK(cudaMalloc(&devQo, sizeof(int) * N));
  CK(cudaMalloc(&devKv, sizeof(int) * N));
  CK(cudaMemcpy(devQo, qoLen.data(), sizeof(int) * N, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(devKv, kvLen.data(), sizeof(int) * N, cudaMemcpyHostToDevice));
  std::vector<int> lo(qoSeq, 0), hi(qoSeq, kvSeq);

  size_t nw = weightSize / sizeof(float);
  std::vector<float> q((size_t)N * qoSeq * qSize), k((size_t)N * kvSeq * kSize),
      v((size_t)N * kvSeq * vSize), w(nw), dout((size_t)N * qoSeq * oProj);
  for (auto &x : q) x = urand();
  for (auto &x : k) x = urand();
  for (auto &x : v) x = urand();
  for (auto &x : w) x = urand() * 0.125f;
  for (auto &x : dout) x = urand();

  float *dq_, *dk_, *dv_, *dw_, *dout_, *out_, *gq, *gk, *gv, *gw;
  void *work, *reserve;
  auto up = [&](float **p, std::vector<float> const &src) {
    CK(cudaMalloc(p, src.size() * sizeof(float)));
    CK(cudaMemcpy(*p, src.data(), src.size() * sizeof(float), cudaMemcpyHostToDevice));
  };
  up(&dq_, q); up(&dk_, k); up(&dv_, v); up(&dw_, w); up(&dout_, dout);
  CK(cudaMalloc(&out_, dout.size() * sizeof(float)));
  CK(cudaMalloc(&gq, q.size() * sizeof(float)));
  CK(cudaMalloc(&gk, k.size() * sizeof(float)));
  CK(cudaMalloc(&gv, v.size() * sizeof(float)));
  CK(cudaMalloc(&gw, weightSize));
  CK(cudaMemset(gw, 0, weightSize));
  CK(cudaMalloc(&work, workSize ? workSize : 16));
  CK(cudaMalloc(&reserve, reserveSize ? reserveSize : 16));

  // ---- weight layout
  std::string lp = dir + "/layout.json";
  FILE *lf = fopen(lp.c_str(), "w");
  fprintf(lf, "{\"weightSizeBytes\": %zu, \"workSpaceBytes\": %zu, \"reserveSpaceBytes\": %zu,\n", weightSize,
          workSize, reserveSize);
  fprintf(lf, " \"cudnn_version\": %zu,\n \"kinds\": {\n", cudnnGetVersion());
  char const *names[4] = {"Q", "K", "V", "O"};
  cudnnMultiHeadAttnWeightKind_t kinds[4] = {CUDNN_MH_ATTN_Q_WEIGHTS, CUDNN_MH_ATTN_K_WEIGHTS,
                                             CUDNN_MH_ATTN_V_WEIGHTS, CUDNN_MH_ATTN_O_WEIGHTS};
  for (int i = 0; i < 4; ++i) {
    cudnnTensorDescriptor_t wd;
    CK(cudnnCreateTensorDescriptor(&wd));
    void *addr = nullptr;
    CK(cudnnGetMultiHeadAttnWeights(h, attn, kinds[i], weightSize, dw_, wd, &addr));
    cudnnDataType_t dt;
    int nb, dims[8], strides[8];
    CK(cudnnGetTensorNdDescriptor(wd, 8, &dt, &nb, dims, strides));
    fprintf(lf, "  \"%s\": {\"offset_elems\": %zu, \"dims\": [", names[i],
            (size_t)((char *)addr - (char *)dw_) / sizeof(float));
    for (int d = 0; d < nb; ++d) fprintf(lf, "%d%s", dims[d], d + 1 < nb ? ", " : "");
    fprintf(lf, "], \"strides\": [");
    for (int d = 0; d < nb; ++d) fprintf(lf, "%d%s", strides[d], d + 1 < nb ? ", " : "");
    fprintf(lf, "]}%s\n", i < 3 ? "," : "");
  }
  fprintf(lf, " },\n \"config\": {\"N\": %d, \"H\": %d, \"qSize\": %d, \"kSize\": %d, \"vSize\": %d, \"qProj\": %d, "
              "\"kProj\": %d, \"vProj\": %d, \"oProj\": %d, \"qoSeq\": %d, \"kvSeq\": %d}}\n",
          N, H, qSize, kSize, vSize, qProj, kProj, vProj, oProj, qoSeq, kvSeq);
  fclose(lf);

  // ---- forward / backward exactly as the reference calls them
  CK(cudnnMultiHeadAttnForward(h, attn, -1, lo.data(), hi.data(), devQo, devKv, qD, dq_, nullptr, kD, dk_,
                               vD, dv_, oD, out_, weightSize, dw_, workSize, work, reserveSize, reserve));
  CK(cudnnMultiHeadAttnBackwardData(h, attn, lo.data(), hi.data(), devQo, devKv, oD, dout_, qD, gq, dq_, kD,
                                    gk, dk_, vD, gv, dv_, weightSize, dw_, workSize, work, reserveSize,
                                    reserve));
  CK(cudnnMultiHeadAttnBackwardWeights(h, attn, CUDNN_WGRAD_MODE_ADD, qD, dq_, kD, dk_, vD, dv_, oD, dout_,
                                       weightSize, dw_, gw, workSize, work, reserveSize, reserve));
  CK(cudaDeviceSynchronize());
  auto down = [&](float *p, size_t n) {
    std::vector<float> r(n);
    CK(cudaMemcpy(r.data(), p, n * sizeof(float), cudaMemcpyDeviceToHost));
    return r;
  };
  dump(dir, "q", q); dump(dir, "k", k); dump(dir, "v", v); dump(dir, "w", w); dump(dir, "dout", dout);
  dump(dir, "out", down(out_, dout.size()));
  dump(dir, "dq", down(gq, q.size()));
  dump(dir, "dk", down(gk, k.size()));
  dump(dir, "dv", down(gv, v.size()));
  dump(dir, "dw", down(gw, nw));
  printf("ok: weightSize=%zu work=%zu reserve=%zu\n", weightSize, workSize, reserveSize);
  return 0;
}
