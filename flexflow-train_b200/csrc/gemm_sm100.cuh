// Dense contraction engine for sm_100a.
//
//   D[b][m][n] (+)= sum_k A(b,m,k) * B(b,n,k)   (+ bias[n]) (activation) (+ lambda * R[m][n])
//
// One warp-specialised persistent kernel: warp 0 drives TMA, warp 1 issues
// tcgen05.mma into a double-buffered TMEM accumulator, warps 2..5 drain TMEM
// with tcgen05.ld and run the fused epilogue. Operands may be K-major
// (contraction index contiguous in memory) or MN-major (row/column index
// contiguous), which covers forward (NT), data-gradient (NN) and
// weight-gradient (TN) products without any transposed copies.
//
// This header is internal to csrc/; the C-ABI lives in include/ffk.h.
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace ffk {
namespace gemm {

enum Act : int { ACT_NONE = -1, ACT_RELU = 0, ACT_SIGMOID = 1, ACT_TANH = 2, ACT_GELU = 3 };

// What to do with an accumulator tile once it is in registers.
struct Epilogue {
  void *d;               // output, row-major [batch][M][ldd]
  long long ldd;         // elements between rows of D
  long long d_batch;     // elements between batches of D
  void const *bias;      // optional, length N, same dtype as D
  int act;               // Act
  int accumulate;        // 1: D += result (beta = 1)
  float alpha;           // scale on the product
  float lambda;          // optional regulariser: D += lambda * R
  void const *r;         // R, same layout/dtype as D (may be null)
  int bias_vec_ok = 0;   // set by the launcher: bias readable as 16-byte packets
  long long d_batch2 = 0;  // elements between inner batches of D (see Problem::batch2)
  int add_d_before_act = 0;  // 1: v = alpha*acc + D, THEN bias / activation (split-precision passes)
  // ---- fused-operator extras (ffk_linear_forward_fused / ffk_linear_backward_fused, dense.cu).
  // Only the staged 16-bit epilogue implements them: 16-bit D with 16-byte aligned rows, N a
  // multiple of 32, one batch; `fused_eligible` checks that before any of them is set. Order:
  //   v = alpha*acc + bias -> aux_out <- v -> v = act(v) -> v = dropout(v) -> R -> (+ D) -> store
  void *aux_out = nullptr;        // copy of the pre-activation, same layout as D
  int r_mode = 0;                 // 0: v += lambda * R;  1: v = R > 0 ? v * lambda : 0 (gate by a ReLU / dropout output)
  int drop_on = 0;                // dropout behind the activation
  unsigned *drop_mask = nullptr;  // 1 keep bit per element of D ([M][N] dense), may be null
  unsigned drop_thresh = 0;       // keep iff the element's 16-bit Philox field >= drop_thresh
  float drop_scale = 1.f;         // 1 / (1 - rate)
  unsigned long long drop_seed = 0, drop_offset = 0;
};

// Problem in "logical" terms. a/b pointers are the stored matrices:
//   K-major  operand: stored [batch][rows = MN][cols = K], ld = elements between rows
//   MN-major operand: stored [batch][rows = K][cols = MN]
struct Problem {
  void const *a;
  void const *b;
  long long lda, ldb;
  long long a_batch, b_batch;  // elements between batches
  int M, N, K, batch;
  int in_dtype;   // FFK_BFLOAT16 / FFK_HALF / FFK_FLOAT (tf32 on the tensor path)
  int out_dtype;  // FFK_BFLOAT16 / FFK_HALF / FFK_FLOAT
  bool a_mn_major, b_mn_major;
  // Optional second (inner) batch level: the batch index b in [0, batch) is split as
  // (b / batch2, b % batch2) with strides (x_batch, x_batch2). Used by attention, where a
  // [sample][seq][head][dim] tensor is batched over (sample, head).
  int batch2 = 1;
  long long a_batch2 = 0, b_batch2 = 0;
  // set when the operands already live in the registered scratch (ffk_set_gemm_workspace), so
  // that the launch does not repack anything into it
  bool scratch_busy = false;
};

// Tensor-core path. Returns FFK_ERR_UNSUPPORTED (without launching) when the
// operands break a TMA constraint (16-byte aligned base and row pitch).
int launch_tc(Problem const &p, Epilogue const &e, cudaStream_t stream);
bool tc_eligible(Problem const &p, Epilogue const &e);
// true when the staged 16-bit epilogue (and therefore Epilogue's fused-operator extras) serves
// every chunk of this problem
bool fused_eligible(Problem const &p, Epilogue const &e);
// D = softmax over each row of (A B^T): 512-column rows, K-major 16-bit operands (the attention
// scores of a 512-token sequence), normalised in TMEM before they are stored.
bool softmax_gemm_eligible(Problem const &p, Epilogue const &e);
int launch_softmax_gemm(Problem const &p, Epilogue const &e, cudaStream_t stream);

// CUDA-core path: any pitch, fp32 accumulate in strict k order per tile. This
// is the fp32-exact path used for FLOAT tensors and the ragged fallback.
int launch_simt(Problem const &p, Epilogue const &e, cudaStream_t stream);

// fp32 operands on the tensor cores at fp32-class accuracy: each operand is split into two
// bf16 terms (x = hi + lo, |x - hi - lo| <= 2^-18 |x|) and the product is formed as
// A_hi B_hi + A_hi B_lo + A_lo B_hi with fp32 accumulation in TMEM (relative error ~1e-5;
// the dropped A_lo B_lo term is 2^-18). Needs split_workspace_bytes(p) of scratch.
size_t split_workspace_bytes(Problem const &p);
int launch_split_bf16(Problem const &p, Epilogue const &e, cudaStream_t stream, void *workspace,
                      size_t workspace_bytes);

// Scratch registered by the caller for GEMM-based entry points (the analogue of
// cublasSetWorkspace); used by AUTO to run large fp32 products through launch_split_bf16.
// One slot per device (the current one); uses on different streams are event-ordered.
int set_workspace(void *ptr, size_t bytes);
// AUTO may run large fp32 products through the bf16 split (current device's slot).
int set_fp32_tensor_math(int allow);
int launch_split_bf16_registered(Problem const &p, Epilogue const &e, cudaStream_t stream);

// K-cut schedule of the CTA-pair kernel (see gemm_sm100.cu `SkWalk`, `sk_begin`): the arithmetic
// alone, for introspection and the CPU tests. kcut_plan returns 1 when an M x N x K 16-bit
// product on `pairs` CTA pairs would be cut (given scratch) and fills plan5 = {whole tiles,
// cut tiles, pieces per cut tile (0 = heads and tails), k-blocks per piece / per head, tails per
// helper pair}; kcut_segments lists pair `pair`'s work as (tile, first k-block, end k-block,
// role, slot, count) and returns the number of segments.
int kcut_plan(int M, int N, int K, int pairs, int *plan5);
// Per-device switch of the K-cut schedule (on by default; needs registered scratch to act).
int set_kcut(int enable);
int kcut_segments(int M, int N, int K, int pairs, int pair, int *six, int max_segments);

// Picks the tensor path for 16-bit inputs when eligible; large fp32 products take the
// split-bf16 tensor path when scratch is registered; everything else runs SIMT.
int launch(Problem const &p, Epilogue const &e, cudaStream_t stream);

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));  // MUFU.TANH, rel. error ~2^-11
  return y;
}
// Activations for 16-bit outputs: hardware tanh / exp are exact to well below bf16/half
// resolution (2^-8 / 2^-11) and keep the epilogue off the critical path.
__device__ __forceinline__ float apply_act_fast(float v, int act) {
  switch (act) {
    case ACT_RELU: return fmaxf(v, 0.f);
    case ACT_SIGMOID: return 0.5f * tanh_fast(0.5f * v) + 0.5f;
    case ACT_TANH: return tanh_fast(v);
    case ACT_GELU: {
      constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
      return v * (0.5f + 0.5f * tanh_fast(v * (C * v * v + B)));
    }
    default: return v;
  }
}

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case ACT_RELU: return fmaxf(v, 0.f);
    case ACT_SIGMOID: return 1.f / (1.f + __expf(-v));
    case ACT_TANH: return tanhf(v);
    case ACT_GELU: {
      // tanh form, as the reference's Linear uses (cuda_helper.cu:111-120)
      constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
      return v * (0.5f + 0.5f * tanhf(v * (C * v * v + B)));
    }
    default: return v;
  }
}

}  // namespace gemm
}  // namespace ffk
