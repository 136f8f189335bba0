// MultiHeadAttention on the GEMM engine: projections, per-(sample, head) score / value
// products as two-level batched GEMMs over the [sample][time][head][dim] projections
// (no head transposes are materialised), row softmax in place, output projection.
#include "gemm_sm100.cuh"

namespace ffk {
namespace {

struct Geo {
  long long N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk;
  long long HQ, HK, HV, Rq, Rk;
  // element offsets inside the packed weight buffer
  long long wq, wk, wv, wo, wtotal;
  // element offsets inside reserve / workspace (each rounded to 16 bytes)
  long long r_qp, r_kp, r_vp, r_p, r_oc, r_total;
  long long w_doc, w_dp, w_dqp, w_dkp, w_dvp, w_total;
};

long long round16(long long elems, size_t es) {
  long long per = 16 / (long long)es;
  return (elems + per - 1) / per * per;
}

int make_geo(ffk_mha_config_t const *c, int dtype, Geo &g) {
  FFK_REQUIRE(c, "mha: null config");
  size_t es = dtype_size(dtype);
  FFK_REQUIRE(dtype == FFK_FLOAT || dtype == FFK_BFLOAT16 || dtype == FFK_HALF,
              "mha: dtype must be FLOAT, HALF or BFLOAT16");
  FFK_REQUIRE(c->num_samples >= 0 && c->num_heads > 0 && c->qSize > 0 && c->kSize > 0 &&
                  c->vSize > 0 && c->qProjSize > 0 && c->kProjSize > 0 && c->vProjSize > 0 &&
                  c->oProjSize > 0 && c->qoSeqLength >= 0 && c->kvSeqLength > 0,
              "mha: bad configuration");
  FFK_REQUIRE(c->qProjSize == c->kProjSize, "mha: qProjSize (%d) must equal kProjSize (%d)",
              c->qProjSize, c->kProjSize);
  g.N = c->num_samples; g.H = c->num_heads; g.qS = c->qSize; g.kS = c->kSize; g.vS = c->vSize;
  g.qP = c->qProjSize; g.kP = c->kProjSize; g.vP = c->vProjSize; g.oP = c->oProjSize;
  g.Sq = c->qoSeqLength; g.Sk = c->kvSeqLength;
  g.HQ = g.H * g.qP; g.HK = g.H * g.kP; g.HV = g.H * g.vP;
  g.Rq = g.N * g.Sq; g.Rk = g.N * g.Sk;
  g.wq = 0; g.wk = g.wq + g.qS * g.HQ; g.wv = g.wk + g.kS * g.HK; g.wo = g.wv + g.vS * g.HV;
  g.wtotal = g.wo + g.HV * g.oP;
  long long o = 0;
  g.r_qp = o; o += round16(g.Rq * g.HQ, es);
  g.r_kp = o; o += round16(g.Rk * g.HK, es);
  g.r_vp = o; o += round16(g.Rk * g.HV, es);
  g.r_p = o;  o += round16(g.N * g.H * g.Sq * g.Sk, es);
  g.r_oc = o; o += round16(g.Rq * g.HV, es);
  g.r_total = o;
  o = 0;
  g.w_doc = o; o += round16(g.Rq * g.HV, es);
  g.w_dp = o;  o += round16(g.N * g.H * g.Sq * g.Sk, es);
  g.w_dqp = o; o += round16(g.Rq * g.HQ, es);
  g.w_dkp = o; o += round16(g.Rk * g.HK, es);
  g.w_dvp = o; o += round16(g.Rk * g.HV, es);
  g.w_total = o;
  return FFK_OK;
}

struct Ctx {
  cudaStream_t s;
  int dtype, path;
  size_t es;
  char *at(void const *base, long long elems) const {
    return const_cast<char *>(reinterpret_cast<char const *>(base)) + elems * (long long)es;
  }
};

// D[M][N] (+)= A B with one or two batch levels. *_mn: operand stored with the M/N index contiguous.
int mm(Ctx const &c, void const *a, long long lda, bool a_mn, void const *b, long long ldb,
       bool b_mn, void *d, long long ldd, long long M, long long N, long long K, bool accumulate,
       long long batch = 1, long long batch2 = 1, long long a_b = 0, long long a_b2 = 0,
       long long b_b = 0, long long b_b2 = 0, long long d_b = 0, long long d_b2 = 0,
       void const *residual = nullptr) {
  if (M == 0 || N == 0 || batch == 0) return FFK_OK;
  gemm::Problem p{a, b, lda, ldb, a_b, b_b, (int)M, (int)N, (int)K, (int)batch, c.dtype, c.dtype,
                  a_mn, b_mn};
  p.batch2 = (int)batch2;
  p.a_batch2 = a_b2;
  p.b_batch2 = b_b2;
  gemm::Epilogue e{d, ldd, d_b, nullptr, gemm::ACT_NONE, accumulate ? 1 : 0, 1.f, 0.f, nullptr};
  e.d_batch2 = d_b2;
  if (residual) {  // D = A B + residual, through the epilogue's `lambda * R` term
    e.r = residual;
    e.lambda = 1.f;
  }
  switch (c.path) {
    case FFK_GEMM_TENSOR: return gemm::launch_tc(p, e, c.s);
    case FFK_GEMM_SIMT: return gemm::launch_simt(p, e, c.s);
    case FFK_GEMM_SPLIT_BF16: return gemm::launch_split_bf16_registered(p, e, c.s);
    default: return gemm::launch(p, e, c.s);
  }
}

}  // namespace
}  // namespace ffk

using namespace ffk;

extern "C" {

size_t ffk_mha_weight_bytes(ffk_mha_config_t const *cfg_host, int dtype) {
  Geo g;
  if (make_geo(cfg_host, dtype, g)) return 0;
  return size_t(g.wtotal) * dtype_size(dtype);
}
size_t ffk_mha_reserve_bytes(ffk_mha_config_t const *cfg_host, int dtype) {
  Geo g;
  if (make_geo(cfg_host, dtype, g)) return 0;
  return size_t(g.r_total) * dtype_size(dtype);
}
size_t ffk_mha_workspace_bytes(ffk_mha_config_t const *cfg_host, int dtype) {
  Geo g;
  if (make_geo(cfg_host, dtype, g)) return 0;
  return size_t(g.w_total) * dtype_size(dtype);
}

}  // extern "C"

namespace {
int mha_forward_impl(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                     void const *key, void const *value, void const *weight, void *output,
                     void *reserve, int dtype, int path, void const *residual) {
  Geo g;
  int rc = make_geo(cfg_host, dtype, g);
  if (rc) return rc;
  if (g.N == 0 || g.Sq == 0) return FFK_OK;
  FFK_REQUIRE(query && key && value && weight && output && reserve, "mha forward: null tensor");
  FFK_REQUIRE(aligned16(reserve), "mha forward: reserve space must be 16-byte aligned");
  Ctx c{as_stream(stream), dtype, path, dtype_size(dtype)};
  void *qp = c.at(reserve, g.r_qp), *kp = c.at(reserve, g.r_kp), *vp = c.at(reserve, g.r_vp);
  void *pp = c.at(reserve, g.r_p), *oc = c.at(reserve, g.r_oc);
  // input projections: X [rows][in] times W [in][H*proj]
  if ((rc = mm(c, query, g.qS, false, c.at(weight, g.wq), g.HQ, true, qp, g.HQ, g.Rq, g.HQ, g.qS, false))) return rc;
  if ((rc = mm(c, key, g.kS, false, c.at(weight, g.wk), g.HK, true, kp, g.HK, g.Rk, g.HK, g.kS, false))) return rc;
  if ((rc = mm(c, value, g.vS, false, c.at(weight, g.wv), g.HV, true, vp, g.HV, g.Rk, g.HV, g.vS, false))) return rc;
  // scores S[n][h] = Qh Kh^T (scale 1.0), batched over (sample, head), then P = softmax rows.
  // 512-key rows of 16-bit tensors are normalised inside the score GEMM (in TMEM); anything
  // else is the GEMM followed by the in-place row softmax.
  {
    gemm::Problem p{qp, kp, g.HQ, g.HK, g.Sq * g.HQ, g.Sk * g.HK, (int)g.Sq, (int)g.Sk, (int)g.qP,
                    (int)(g.N * g.H), dtype, dtype, false, false};
    p.batch2 = (int)g.H;
    p.a_batch2 = g.qP;
    p.b_batch2 = g.kP;
    gemm::Epilogue e{pp, g.Sk, g.H * g.Sq * g.Sk, nullptr, gemm::ACT_NONE, 0, 1.f, 0.f, nullptr};
    e.d_batch2 = g.Sq * g.Sk;
    bool const tensor_path = path == FFK_GEMM_AUTO || path == FFK_GEMM_TENSOR;
    if (tensor_path && gemm::softmax_gemm_eligible(p, e)) {
      if ((rc = gemm::launch_softmax_gemm(p, e, c.s))) return rc;
    } else {
      if ((rc = mm(c, qp, g.HQ, false, kp, g.HK, false, pp, g.Sk, g.Sq, g.Sk, g.qP, false, g.N * g.H,
                   g.H, g.Sq * g.HQ, g.qP, g.Sk * g.HK, g.kP, g.H * g.Sq * g.Sk, g.Sq * g.Sk)))
        return rc;
      if ((rc = ffk_softmax_forward(stream, pp, pp, g.N * g.H * g.Sq, g.Sk, 1, dtype))) return rc;
    }
  }
  // head outputs O[n][:, h, :] = P[n][h] Vh, written straight into the concatenated layout
  if ((rc = mm(c, pp, g.Sk, false, vp, g.HV, true, oc, g.HV, g.Sq, g.vP, g.Sk, false, g.N * g.H, g.H,
               g.H * g.Sq * g.Sk, g.Sq * g.Sk, g.Sk * g.HV, g.vP, g.Sq * g.HV, g.vP)))
    return rc;
  // output projection (+ residual)
  return mm(c, oc, g.HV, false, c.at(weight, g.wo), g.oP, true, output, g.oP, g.Rq, g.oP, g.HV, false,
            1, 1, 0, 0, 0, 0, 0, 0, residual);
}

// acc_*: 0 = the gradient is overwritten (cuDNN's BackwardData semantics), 1 = added to.
int mha_backward_impl(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                      void *query_grad, void const *key, void *key_grad, void const *value,
                      void *value_grad, void const *weight, void *weight_grad,
                      void const *output_grad, void const *reserve, void *workspace,
                      size_t workspace_bytes, int dtype, int path, bool acc_q, bool acc_k, bool acc_v) {
  Geo g;
  int rc = make_geo(cfg_host, dtype, g);
  if (rc) return rc;
  if (g.N == 0 || g.Sq == 0) return FFK_OK;
  FFK_REQUIRE(query && key && value && weight && weight_grad && output_grad && reserve && workspace,
              "mha backward: null tensor");
  FFK_REQUIRE(workspace_bytes >= size_t(g.w_total) * dtype_size(dtype),
              "mha backward: workspace of %zu bytes is smaller than the %zu required",
              workspace_bytes, size_t(g.w_total) * dtype_size(dtype));
  FFK_REQUIRE(aligned16(reserve) && aligned16(workspace), "mha backward: unaligned scratch");
  Ctx c{as_stream(stream), dtype, path, dtype_size(dtype)};
  void *qp = c.at(reserve, g.r_qp), *kp = c.at(reserve, g.r_kp), *vp = c.at(reserve, g.r_vp);
  void *pp = c.at(reserve, g.r_p), *oc = c.at(reserve, g.r_oc);
  void *doc = c.at(workspace, g.w_doc), *dp = c.at(workspace, g.w_dp);
  void *dqp = c.at(workspace, g.w_dqp), *dkp = c.at(workspace, g.w_dkp), *dvp = c.at(workspace, g.w_dvp);
  void const *wq = c.at(weight, g.wq), *wk = c.at(weight, g.wk), *wv = c.at(weight, g.wv), *wo = c.at(weight, g.wo);
  long long const NH = g.N * g.H;
  // output projection: dOc = dOut Wo^T ; dWo += Oc^T dOut
  if ((rc = mm(c, output_grad, g.oP, false, wo, g.oP, false, doc, g.HV, g.Rq, g.HV, g.oP, false))) return rc;
  if ((rc = mm(c, oc, g.HV, true, output_grad, g.oP, true, c.at(weight_grad, g.wo), g.oP, g.HV, g.oP, g.Rq, true))) return rc;
  // dP = dOc_h Vh^T ; dVp_h = P^T dOc_h
  if ((rc = mm(c, doc, g.HV, false, vp, g.HV, false, dp, g.Sk, g.Sq, g.Sk, g.vP, false, NH, g.H,
               g.Sq * g.HV, g.vP, g.Sk * g.HV, g.vP, g.H * g.Sq * g.Sk, g.Sq * g.Sk)))
    return rc;
  if ((rc = mm(c, pp, g.Sk, true, doc, g.HV, true, dvp, g.HV, g.Sk, g.vP, g.Sq, false, NH, g.H,
               g.H * g.Sq * g.Sk, g.Sq * g.Sk, g.Sq * g.HV, g.vP, g.Sk * g.HV, g.vP)))
    return rc;
  // dS = P (.) (dP - rowsum(dP (.) P)), in place on dP
  if ((rc = ffk_softmax_backward(stream, dp, dp, pp, NH * g.Sq, g.Sk, 1, dtype, 1))) return rc;
  // dQp_h = dS Kh ; dKp_h = dS^T Qh
  if ((rc = mm(c, dp, g.Sk, false, kp, g.HK, true, dqp, g.HQ, g.Sq, g.qP, g.Sk, false, NH, g.H,
               g.H * g.Sq * g.Sk, g.Sq * g.Sk, g.Sk * g.HK, g.kP, g.Sq * g.HQ, g.qP)))
    return rc;
  if ((rc = mm(c, dp, g.Sk, true, qp, g.HQ, true, dkp, g.HK, g.Sk, g.kP, g.Sq, false, NH, g.H,
               g.H * g.Sq * g.Sk, g.Sq * g.Sk, g.Sq * g.HQ, g.qP, g.Sk * g.HK, g.kP)))
    return rc;
  // input projections: dX = dXp W^T (overwrite) ; dW += X^T dXp
  if (query_grad && (rc = mm(c, dqp, g.HQ, false, wq, g.HQ, false, query_grad, g.qS, g.Rq, g.qS, g.HQ, acc_q))) return rc;
  if (key_grad && (rc = mm(c, dkp, g.HK, false, wk, g.HK, false, key_grad, g.kS, g.Rk, g.kS, g.HK, acc_k))) return rc;
  if (value_grad && (rc = mm(c, dvp, g.HV, false, wv, g.HV, false, value_grad, g.vS, g.Rk, g.vS, g.HV, acc_v))) return rc;
  if ((rc = mm(c, query, g.qS, true, dqp, g.HQ, true, c.at(weight_grad, g.wq), g.HQ, g.qS, g.HQ, g.Rq, true))) return rc;
  if ((rc = mm(c, key, g.kS, true, dkp, g.HK, true, c.at(weight_grad, g.wk), g.HK, g.kS, g.HK, g.Rk, true))) return rc;
  return mm(c, value, g.vS, true, dvp, g.HV, true, c.at(weight_grad, g.wv), g.HV, g.vS, g.HV, g.Rk, true);
}
}  // namespace

extern "C" {

int ffk_mha_forward(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                    void const *key, void const *value, void const *weight, void *output,
                    void *reserve, int dtype, int path) {
  return mha_forward_impl(stream, cfg_host, query, key, value, weight, output, reserve, dtype, path, nullptr);
}

int ffk_mha_forward_residual(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                             void const *key, void const *value, void const *weight, void *output,
                             void *reserve, int dtype, int path, void const *residual) {
  return mha_forward_impl(stream, cfg_host, query, key, value, weight, output, reserve, dtype, path, residual);
}

int ffk_mha_backward(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                     void *query_grad, void const *key, void *key_grad, void const *value,
                     void *value_grad, void const *weight, void *weight_grad,
                     void const *output_grad, void const *reserve, void *workspace,
                     size_t workspace_bytes, int dtype, int path) {
  return mha_backward_impl(stream, cfg_host, query, query_grad, key, key_grad, value, value_grad, weight,
                           weight_grad, output_grad, reserve, workspace, workspace_bytes, dtype, path, false,
                           false, false);
}

int ffk_mha_backward_accumulate(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                                void *query_grad, void const *key, void *key_grad, void const *value,
                                void *value_grad, void const *weight, void *weight_grad,
                                void const *output_grad, void const *reserve, void *workspace,
                                size_t workspace_bytes, int dtype, int path, int accumulate_query_grad,
                                int accumulate_key_grad, int accumulate_value_grad) {
  return mha_backward_impl(stream, cfg_host, query, query_grad, key, key_grad, value, value_grad, weight,
                           weight_grad, output_grad, reserve, workspace, workspace_bytes, dtype, path,
                           accumulate_query_grad != 0, accumulate_key_grad != 0, accumulate_value_grad != 0);
}

}  // extern "C"
