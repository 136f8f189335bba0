/* ffk_dense.h — tensor-core bound entry points: Linear, BatchMatmul and the GEMM they share.
 * Included by ffk.h. */
#ifndef FFK_DENSE_H_
#define FFK_DENSE_H_
#include "ffk.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Which contraction engine a call uses. AUTO = tcgen05/TMEM/TMA tiles for 16-bit inputs
 * whose pitches are 16-byte aligned, CUDA-core fp32 (exact fp32 products, fp32
 * accumulation) otherwise. TENSOR on FLOAT inputs runs them as TF32 (10-bit mantissa). */
typedef enum { FFK_GEMM_AUTO = 0, FFK_GEMM_TENSOR = 1, FFK_GEMM_SIMT = 2, FFK_GEMM_SPLIT_BF16 = 3 } ffk_gemm_path_t;

/* Registers device scratch for the GEMM-based entry points (Linear, BatchMatmul,
 * MultiHeadAttention, Conv2D, ffk_gemm) ON THE CURRENT DEVICE — the role cublasSetWorkspace /
 * PerDeviceFFHandle::workSpace (include/kernels/ff_handle.h:17-18) play for the reference's
 * cuBLAS calls. One slot per device: a process that drives several GPUs registers one buffer
 * per device, with that device current; the pointer must be memory of that device. With
 * scratch registered, 16-bit operands that break TMA's 16-byte pitch rule are repacked into it
 * and stay on the tensor cores. Calls on different streams of one device that need the scratch
 * are ordered behind each other with an event (inside a stream capture that is impossible and
 * the call takes the path without scratch). NULL unregisters. */
int ffk_set_gemm_workspace(void *workspace, size_t workspace_bytes);

/* PerDeviceFFHandle::allowTensorOpMathConversion (include/kernels/ff_handle.h:19; the reference's
 * handle sets it to true, src/managed_per_device_ff_handle.cc:9). allow != 0: FFK_GEMM_AUTO runs
 * large fp32 products of the current device on the tensor cores by splitting each fp32 operand
 * into two bf16 terms and summing three tcgen05 products in fp32 (relative error ~1e-5, inside
 * the fp32 tolerance of 1e-4) — needs registered scratch of 2 bytes x 2 x (elements of A + B) per
 * call. allow == 0 (the default): fp32 products stay exact on the CUDA cores whatever is
 * registered. FFK_GEMM_SPLIT_BF16 as an explicit path needs only the scratch. */
int ffk_set_gemm_fp32_tensor_math(int allow);

/* The K-cut schedule of the 16-bit tensor-core product, as arithmetic (no device work; no
 * reference counterpart — cuBLAS picks its own split-K heuristics behind cublasGemmEx,
 * linear_kernels.cu:127-149). A product of more than 128 x 128 outputs runs as 256 x 256 tiles
 * on `pairs` CTA pairs (SM count / 2). When the tile count leaves >= 5 % of the launch idle in
 * its last wave AND scratch is registered (ffk_set_gemm_workspace, >= 16 KB + pairs x 256 KB),
 * the tiles of that wave are cut along K (k-blocks of 64): into equal pieces, one pair each,
 * when the wave is at most half full; otherwise into a head computed by the tile's own pair and
 * a tail, with the idle pairs taking several tails each. Partial sums are parked in the scratch
 * in fp32 and added by the tile's owner in a fixed order (results are reproducible run to run;
 * they differ from the uncut schedule by fp32 summation order only). FFK_GEMM_NO_STREAMK=1
 * disables it.
 *   ffk_gemm_kcut_plan     -> 1 if M x N x K would be cut; plan5 = {whole tiles, cut tiles,
 *                             pieces per cut tile (0 = heads and tails), k-blocks per piece or
 *                             per head, tails per helper pair} (all 0 when not cut);
 *   ffk_gemm_kcut_segments -> pair `pair`'s work as (tile, first k-block, end k-block, role,
 *                             slot, count) written to six[6 * max_segments]; role 0 = whole
 *                             tile, 1 = partial parked in `slot`, 2 = owner adding `count`
 *                             partials from `slot` on; returns the segment count. */
int ffk_gemm_kcut_plan(int M, int N, int K, int pairs, int *plan5);
/* Per-device switch of the K-cut schedule for the CURRENT device (default on; it acts only with
 * registered scratch). Off = whole tiles only — for callers that run kernels which wait on their
 * own grid's CTAs beside 16-bit products at equal stream priority ("Residency").
 * No reference counterpart (cuBLAS exposes no such switch; linear_kernels.cu:127-149). */
int ffk_set_gemm_kcut(int enable);
int ffk_gemm_kcut_segments(int M, int N, int K, int pairs, int pair, int *six, int max_segments);

/* General strided-batched product, the building block under Linear / BatchMatmul / MHA:
 *   D[b][m][n] = alpha * sum_k A(b,m,k) * B(b,n,k)  (+ bias[n]) (act)  (+ D if accumulate)
 * A is stored [batch][M][lda>=K] when a_mn_major == 0 (k contiguous) or [batch][K][lda>=M]
 * when a_mn_major == 1 (m contiguous); likewise B with N. D is [batch][M][ldd>=N].
 * Replaces the cublasGemmEx / cublasSgemmStridedBatched calls at
 * src/cuda/ops/linear_kernels.cu:131,231,303 and src/cuda/ops/batch_matmul_kernels.cu:66,105,123. */
int ffk_gemm(void *stream, void const *a, void const *b, void *d, int M, int N, int K, int batch,
             long long lda, long long ldb, long long ldd, long long a_batch_stride,
             long long b_batch_stride, long long d_batch_stride, int in_dtype, int out_dtype,
             int a_mn_major, int b_mn_major, float alpha, int accumulate, void const *bias,
             int activation, int path);

/* Kernels::Linear::forward_kernel — include/kernels/linear_kernels.h:52,
 * src/cuda/ops/linear_kernels.cu:109-190.
 *   Y[b][o] = act( sum_i X[b][i] * W[o][i] + bias[o] ),  W stored [out_dim][in_dim].
 * bias may be NULL. GELU is the tanh form the reference's Linear uses. Accumulation is fp32
 * (the reference asks cuBLAS for COMPUTE_16F; see DESIGN.md "deviations"). */
int ffk_linear_forward(void *stream, void const *input, void *output, void const *weight,
                       void const *bias, int in_dim, int out_dim, int batch_size, int dtype,
                       int activation, int path);

/* Kernels::Linear::backward_kernel — include/kernels/linear_kernels.h:62,
 * src/cuda/ops/linear_kernels.cu:192-323. In this order, as the reference does:
 *   dY <- dY (.) act'(Y) in place (RELU, SIGMOID as the reference; TANH and tanh-GELU added);
 *   dW[o][i] += sum_b dY[b][o] X[b][i]  (+ l2_lambda * W[o][i] when l2_lambda != 0);
 *   db[o]   += sum_b dY[b][o]           (bias_grad may be NULL);
 *   dX[b][i] += sum_o dY[b][o] W[o][i]  (input_grad may be NULL).
 * The activation gradient is evaluated from the post-activation `output` (Y) for
 * RELU / SIGMOID / TANH. GELU needs the pre-activation values: pass them in `preact`
 * (same shape as Y); with preact == NULL a GELU request fails with FFK_ERR_UNSUPPORTED
 * (the reference asserts for anything but RELU / SIGMOID).
 * `workspace` (device scratch, optional — the reference keeps one in
 * PerDeviceFFHandle::workSpace, include/kernels/ff_handle.h:17-18) lets the bias gradient run
 * as a two-phase column reduction at HBM speed: ffk_linear_backward_workspace_bytes(out_dim)
 * bytes; with NULL a single-pass kernel without scratch is used. Its contents are not
 * preserved between calls. */
size_t ffk_linear_backward_workspace_bytes(int out_dim);
int ffk_linear_backward(void *stream, void const *input, void *input_grad, void const *output,
                        void *output_grad, void const *weight, void *weight_grad,
                        void *bias_grad, int in_dim, int out_dim, int batch_size, int dtype,
                        int activation, float l2_lambda, void const *preact, int path,
                        void *workspace, size_t workspace_bytes);

/* ---- Fused-operator execution (the reference's `perform_fusion` switch,
 * lib/local-execution/include/local-execution/config.h:80, and the legacy FusedOp,
 * lib/runtime/src/ops/fused.cc: consecutive operators run as one task). These entry points are
 * ADDITIVE: each equals a chain of the per-operator entry points above and falls back to exactly
 * that chain when a shape or dtype cannot take the one-kernel path (16-bit tensors, out_dim a
 * multiple of 32, 16-byte aligned rows). */

/* Linear -> [activation] -> [Dropout] -> [+ residual], the feed-forward and projection blocks of
 * lib/models (transformer.cc:46-60, bert.cc:98-120), as ONE GEMM whose epilogue does the rest:
 *   v = X W^T + bias;  preact_out <- v (optional; the GELU gradient needs it);  v = act(v);
 *   v = keep ? v / (1 - rate) : 0 with the keep bit of ffk_dropout_forward's generator for the
 *   same (seed, offset, element) — the bit mask is also written when dropout_mask != NULL;
 *   output <- v + residual (optional).
 * Replaces linear_kernels.cu:109-190 + dropout_kernels.cu:64-80 + element_binary_kernels.cu:141-243. */
typedef struct {
  void *preact_out;            /* optional, same shape and dtype as output */
  void const *residual;        /* optional, same shape and dtype as output */
  void *dropout_mask;          /* ffk_dropout_mask_bytes(batch*out_dim) bytes; needed by the backward pass unless the activation is RELU */
  float dropout_rate;          /* 0 = no dropout */
  unsigned long long dropout_seed, dropout_offset;
} ffk_linear_forward_fusion_t;
int ffk_linear_forward_fused(void *stream, void const *input, void *output, void const *weight,
                             void const *bias, int in_dim, int out_dim, int batch_size, int dtype,
                             int activation, int path, ffk_linear_forward_fusion_t const *fusion_host);

/* ffk_linear_backward (linear_kernels.h:62, linear_kernels.cu:192-323) with the operators around it
 * folded in (the dropout gradient of dropout_kernels.cu:82-94, the ReLU gradient of the layer before):
 *   accumulate_input_grad   1: dX += (the reference's convention); 0: dX = — the caller knows this
 *                           Linear is the first writer of that gradient, so no zero fill is needed;
 *   dropout_mask / rate     a dropout followed the activation (fused forward above): its gradient
 *                           and the activation gradient are applied to dY in one pass;
 *   output_grad_is_preactivation   dY already holds the pre-activation gradient (it was produced by
 *                           a gated data-gradient epilogue, below): skip that pass;
 *   input_gate / scale      X is the stored output of ReLU (+ dropout, scale = 1/(1-rate)) of the
 *                           layer before: the data-gradient GEMM's epilogue writes
 *                           dX = X > 0 ? (dY W) * scale : 0, i.e. the pre-activation gradient of that
 *                           layer, without a separate pass over the [batch][in_dim] tensor. */
typedef struct {
  int accumulate_input_grad;
  int output_grad_is_preactivation;
  void const *dropout_mask;
  float dropout_rate;
  void const *input_gate;
  float input_gate_scale;
} ffk_linear_backward_fusion_t;
int ffk_linear_backward_fused(void *stream, void const *input, void *input_grad, void const *output,
                              void *output_grad, void const *weight, void *weight_grad,
                              void *bias_grad, int in_dim, int out_dim, int batch_size, int dtype,
                              int activation, float l2_lambda, void const *preact, int path,
                              void *workspace, size_t workspace_bytes,
                              ffk_linear_backward_fusion_t const *fusion_host);

/* Kernels::BatchMatmul::forward_kernel — include/kernels/batch_matmul_kernels.h:12,
 * src/cuda/ops/batch_matmul_kernels.cu:23-84 (argument order of the DEFINITION).
 *   O[b] (n x m) = A[b] (n x k) * B[b] (k x m), all row-major; a/b_seq_length_dim and
 *   seq_length truncate k / n / m exactly as the reference does (negative = off). */
int ffk_batch_matmul_forward(void *stream, void *output, void const *a_input, void const *b_input,
                             int m, int n, int k, int batch, int a_seq_length_dim,
                             int b_seq_length_dim, int seq_length, int dtype, int path);

/* Kernels::BatchMatmul::backward_kernel — include/kernels/batch_matmul_kernels.h:25,
 * src/cuda/ops/batch_matmul_kernels.cu:86-141.
 *   dA[b] += dO[b] * B[b]^T ;  dB[b] += A[b]^T * dO[b]. `o` is unused, as in the reference. */
int ffk_batch_matmul_backward(void *stream, void const *o, void const *o_grad, void const *a,
                              void *a_grad, void const *b, void *b_grad, int m, int n, int k,
                              int batch, int dtype, int path);

/* Kernels::MultiHeadAttention — include/kernels/attention_kernels.h:70-106,
 * src/cuda/ops/attention_kernels.cu:59-345 (cuDNN legacy multi-head attention: ALL_TO_ONE,
 * smScalar 1.0 i.e. NO 1/sqrt(d) scaling, no biases, no dropout, full attention window,
 * sequence data [batch][time][vect]). The fields are init_kernel's integer arguments. */
typedef struct {
  int num_samples, num_heads;
  int qSize, kSize, vSize;
  int qProjSize, kProjSize, vProjSize, oProjSize;
  int qoSeqLength, kvSeqLength;
} ffk_mha_config_t;

/* Bytes of the packed weight buffer: what cudnnGetMultiHeadAttnBuffers reports as weightSize
 * (attention_kernels.cu:127) for the same configuration, with the same layout —
 *   Wq [qSize][heads][qProj] | Wk [kSize][heads][kProj] | Wv [vSize][heads][vProj] | Wo [heads][vProj][oProj]
 * — verified against cudnnGetMultiHeadAttnWeights on a B200 (tests/golden/mha_*.npz). */
size_t ffk_mha_weight_bytes(ffk_mha_config_t const *cfg_host, int dtype);
/* Bytes the forward pass keeps for the backward pass (the role of MHAPerDeviceState::
 * reserveSpace, sized by cudnnGetMultiHeadAttnBuffers at attention_kernels.cu:127-136): projected
 * Q/K/V, attention probabilities, concatenated head outputs. */
size_t ffk_mha_reserve_bytes(ffk_mha_config_t const *cfg_host, int dtype);
/* Scratch bytes backward needs (the role of PerDeviceFFHandle::workSpace, include/kernels/ff_handle.h:13-24,
 * which attention_kernels.cu:292-334 hands to cuDNN). */
size_t ffk_mha_workspace_bytes(ffk_mha_config_t const *cfg_host, int dtype);

/* MultiHeadAttention::forward_kernel — attention_kernels.h:85, attention_kernels.cu:246-277.
 * out[n][s][:] = concat_h( softmax(Qh Kh^T) Vh ) Wo with Qh = q Wq_h etc. `reserve` must hold
 * ffk_mha_reserve_bytes() and is read by the backward call. */
int ffk_mha_forward(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                    void const *key, void const *value, void const *weight, void *output,
                    void *reserve, int dtype, int path);

/* MultiHeadAttention::backward_kernel — attention_kernels.h:93, attention_kernels.cu:279-334.
 * query/key/value gradients are OVERWRITTEN (cuDNN's BackwardData semantics; key_grad and
 * value_grad may be NULL, which the reference's wrapper passes for self-attention,
 * lib/local-execution/src/ops/attention.cc:177-181); weight_grad accumulates
 * (CUDNN_WGRAD_MODE_ADD). */
int ffk_mha_backward(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                     void *query_grad, void const *key, void *key_grad, void const *value,
                     void *value_grad, void const *weight, void *weight_grad,
                     void const *output_grad, void const *reserve, void *workspace,
                     size_t workspace_bytes, int dtype, int path);

/* Fused-operator forms of MultiHeadAttention (additive, see "Fused-operator execution" above):
 *  - forward with the residual of the encoder layer folded into the output projection's epilogue:
 *    out = MHA(q, k, v) + residual (lib/models transformer.cc:32-45 adds the layer input right
 *    behind the attention block);
 *  - backward whose data gradients are ADDED to query_grad / key_grad / value_grad where the flag
 *    says so, instead of cuDNN's overwrite. With q = k = v = x (self-attention) the three pointers
 *    may be the same buffer: accumulate flags (first-writer?, 1, 1) leave dq + dk + dv in it
 *    without three separate tensors and a sum pass. */
int ffk_mha_forward_residual(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                             void const *key, void const *value, void const *weight, void *output,
                             void *reserve, int dtype, int path, void const *residual);
int ffk_mha_backward_accumulate(void *stream, ffk_mha_config_t const *cfg_host, void const *query,
                                void *query_grad, void const *key, void *key_grad, void const *value,
                                void *value_grad, void const *weight, void *weight_grad,
                                void const *output_grad, void const *reserve, void *workspace,
                                size_t workspace_bytes, int dtype, int path, int accumulate_query_grad,
                                int accumulate_key_grad, int accumulate_value_grad);

#ifdef __cplusplus
}
#endif
#endif
