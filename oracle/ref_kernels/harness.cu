// oracle/_ref/libref_kernels.so — the reference's OWN hand-written CUDA kernels, compiled from
// the reference's sources (cut out at build time by extract.py into oracle/_ref/*.inc; nothing of
// them is tracked here), behind a small C interface. Test infrastructure: it gives the oracle and
// the sm_100a kernels a second referee that IS the reference's code, for every op family whose
// arithmetic the reference wrote itself (LayerNorm, Embedding, losses, optimizers, Gather, Cast,
// TopK, the element-unary custom ops, the helper kernels, metrics, Replicate / Reduction sums),
// plus the cuBLAS calls of Linear and BatchMatmul (forward and backward) configured as the reference
// configures them.
//
// Each wrapper restates the launch configuration of the reference's host function (cited), which
// cannot be compiled here (generated *.dtg.h headers, cuDNN / cuBLAS handle types, Legion).
// All pointers are device pointers; everything runs on the given stream.
#include <cublas_v2.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <optional>

// ---- what the cut-out device code expects from the reference's headers
// lib/kernels/src/device.h:59-71, include/kernels/device.h:86
#define CUDA_KERNEL_LOOP(i, n) \
  for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += blockDim.x * gridDim.x)
#define MAX_TENSOR_DIM 5

namespace FlexFlow {
using coord_t = long long;
int const CUDA_NUM_THREADS = 1024;
int const BLOCK_SIZE_LIMIT = 32768;
inline int GET_BLOCKS(int const N) {
  int ret = (N + CUDA_NUM_THREADS - 1) / CUDA_NUM_THREADS;
  return (ret > BLOCK_SIZE_LIMIT) ? BLOCK_SIZE_LIMIT : ret;
}
// op-attrs enums the kernels switch on (only the members they name; values are private to this file)
enum class OperatorType {
  EXP, IDENTITY, RELU, SIGMOID, TANH, ELU, GELU, RSQRT, SIN, COS,
  SCALAR_MULTIPLY, SCALAR_ADD, SCALAR_SUB, SCALAR_TRUE_DIV, POW
};
enum class AggregateOp { SUM, AVG };
// legacy metrics structs (lib/runtime/include/runtime/metrics_functions.h of the legacy tree)
struct PerfMetrics {
  int train_all, train_correct;
  float cce_loss, sparse_cce_loss, mse_loss, rmse_loss, mae_loss;
};
struct Metrics {
  bool measure_accuracy, measure_categorical_crossentropy, measure_sparse_categorical_crossentropy,
      measure_mean_squared_error, measure_root_mean_squared_error, measure_mean_absolute_error;
};

namespace ref_ln {
#include "../_ref/layer_norm.inc"
}
namespace ref_emb {
#include "../_ref/embedding.inc"
}
namespace ref_loss {
#include "../_ref/loss.inc"
}
namespace ref_opt {
#include "../_ref/optimizer.inc"
}
namespace ref_gather {
#include "../_ref/gather.inc"
}
namespace ref_transpose {
#include "../_ref/transpose.inc"
}
namespace ref_cast {
#include "../_ref/cast.inc"
}
namespace ref_helper {
#include "../_ref/helper.inc"
}
namespace ref_unary {
#include "../_ref/element_unary.inc"
}
namespace ref_metrics {
#include "../_ref/metrics.inc"
}
namespace ref_topk {
#include "../_ref/topk.inc"
}
namespace ref_replicate {
#include "../_ref/replicate.inc"
}
namespace ref_reduction {
#include "../_ref/reduction.inc"
}
}  // namespace FlexFlow

using namespace FlexFlow;

#define REF_API extern "C" __attribute__((visibility("default")))
static inline cudaStream_t S(void *s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int done() { return cudaGetLastError() == cudaSuccess ? 0 : -1; }

// ------------------------------------------------------------------ LayerNorm
// layer_norm_kernels.cu:325-350 (ForwardKernel): one block of 512 threads per row for the
// moments, one block of 256 threads per row for the affine map.
REF_API int ref_layer_norm_forward(void *stream, float const *x, float const *gamma, float const *beta,
                                   float *y, float *mean, float *rstd, long long M, long long N, float eps) {
  using namespace ref_ln;
  RowwiseMomentsCUDAKernel<float><<<M, kCUDABlockReduceNumThreads, 0, S(stream)>>>(N, eps, x, mean, rstd);
  LayerNormForwardCUDAKernel<float><<<M, kCUDANumThreads, 0, S(stream)>>>(N, x, mean, rstd, gamma, beta, y);
  return done();
}
// layer_norm_kernels.cu:352-408 (BackwardKernel): ds/db per row, fused scale/bias per row, then
// the column reductions for dgamma / dbeta (simple kernel for M < 512, tiled otherwise).
// dX: LayerNormBackwardCUDAKenrel (:163-181) is defined by the reference but never launched;
// launched here (one block of 256 threads per row, like the forward map) when dx != NULL so
// that the dX formula itself is pinned.
REF_API int ref_layer_norm_backward(void *stream, float const *dy, float const *x, float const *gamma,
                                    float const *mean, float const *rstd, float *ds, float *db, float *scale,
                                    float *bias, float *dx, float *dgamma, float *dbeta, long long M, long long N) {
  using namespace ref_ln;
  ComputeInternalGradientsCUDAKernel<float><<<M, kCUDABlockReduceNumThreads, 0, S(stream)>>>(N, dy, x, gamma, ds, db);
  const int64_t B = (M + kCUDANumThreads - 1) / kCUDANumThreads;
  ComputeGradientFusedParamsCUDAKernel<float><<<B, kCUDANumThreads, 0, S(stream)>>>(M, N, mean, rstd, ds, db, scale, bias);
  if (dx) LayerNormBackwardCUDAKenrel<float><<<M, kCUDANumThreads, 0, S(stream)>>>(N, dy, x, gamma, rstd, scale, bias, dx);
  if (M < 512) {
    const int64_t Bc = (N + kCUDANumThreads - 1) / kCUDANumThreads;
    GammaBetaBackwardSimpleCUDAKernel<float><<<Bc, kCUDANumThreads, 0, S(stream)>>>(M, N, dy, x, mean, rstd, dgamma, dbeta);
  } else {
    const int64_t Bc = (N + kColwiseReduceTileSize - 1) / kColwiseReduceTileSize;
    constexpr int kThreadX = kColwiseReduceTileSize;
    constexpr int kThreadY = kColwiseReduceTileSize / 2;
    GammaBetaBackwardCUDAKernel<float><<<Bc, dim3(kThreadX, kThreadY), 0, S(stream)>>>(M, N, dy, x, mean, rstd, dgamma, dbeta);
  }
  return done();
}

// ------------------------------------------------------------------ Embedding
// embedding_kernels.cu:344-426. aggr: -1 none, 0 SUM, 1 AVG. idx64: 1 = int64 indices. half_rows:
// 1 = __half table. The definitions are templates over a NON-TYPE first parameter (:63-335), so
// they are named with a dummy value; the pointer type selects the int32 / int64 overload.
template <typename TI, typename TD>
static int emb_fwd(cudaStream_t s, TI const *in, TD *out, TD const *w, int aggr, int in_dim, int out_dim, int batch) {
  int vol = batch * out_dim;
  if (aggr < 0)
    ref_emb::embed_forward_no_aggr<0, TD><<<GET_BLOCKS(vol), CUDA_NUM_THREADS, 0, s>>>(in, out, w, out_dim, batch);
  else
    ref_emb::embed_forward_with_aggr<0, TD><<<GET_BLOCKS(vol), CUDA_NUM_THREADS, 0, s>>>(
        in, out, w, out_dim, in_dim, batch, std::optional<AggregateOp>(aggr == 0 ? AggregateOp::SUM : AggregateOp::AVG));
  return done();
}
template <typename TI, typename TD>
static int emb_bwd(cudaStream_t s, TI const *in, TD const *dout, TD *dw, int aggr, int in_dim, int out_dim, int batch) {
  int vol = batch * out_dim;
  if (aggr < 0)
    ref_emb::embed_backward_no_aggr<0, TD><<<GET_BLOCKS(vol), CUDA_NUM_THREADS, 0, s>>>(in, dout, dw, out_dim, batch);
  else
    ref_emb::embed_backward_with_aggr<0, TD><<<GET_BLOCKS(vol), CUDA_NUM_THREADS, 0, s>>>(
        in, dout, dw, out_dim, in_dim, batch, std::optional<AggregateOp>(aggr == 0 ? AggregateOp::SUM : AggregateOp::AVG));
  return done();
}
REF_API int ref_embedding_forward(void *stream, void const *idx, void *out, void const *table, int aggr, int in_dim,
                                  int out_dim, int batch, int idx64) {
  return idx64 ? emb_fwd(S(stream), (int64_t const *)idx, (float *)out, (float const *)table, aggr, in_dim, out_dim, batch)
               : emb_fwd(S(stream), (int32_t const *)idx, (float *)out, (float const *)table, aggr, in_dim, out_dim, batch);
}
REF_API int ref_embedding_backward(void *stream, void const *idx, void const *dout, void *dtable, int aggr, int in_dim,
                                   int out_dim, int batch, int idx64) {
  return idx64 ? emb_bwd(S(stream), (int64_t const *)idx, (float const *)dout, (float *)dtable, aggr, in_dim, out_dim, batch)
               : emb_bwd(S(stream), (int32_t const *)idx, (float const *)dout, (float *)dtable, aggr, in_dim, out_dim, batch);
}

// ------------------------------------------------------------------ losses
// loss_function_kernels.cu:59-139: copy / subtract, then scale_kernel(ptr, n, 0, scale) of
// cuda_helper.cu:32-36 (ptr = (b - a) * ptr + a with a = 0).
REF_API int ref_sparse_cce_loss_backward(void *stream, float *g, float const *p, int const *label, int num_samples,
                                         int num_classes, int k, float scale) {
  size_t vol = (size_t)num_samples * num_classes;
  cudaMemcpyAsync(g, p, vol * sizeof(float), cudaMemcpyDeviceToDevice, S(stream));
  ref_loss::sparse_categorical_crossentropy_loss_backward<<<GET_BLOCKS(num_samples), CUDA_NUM_THREADS, 0, S(stream)>>>(
      g, label, num_samples, num_classes, k);
  ref_helper::scale_kernel<<<GET_BLOCKS(vol), CUDA_NUM_THREADS, 0, S(stream)>>>(g, vol, 0, scale * k);
  return done();
}
// kind: 0 categorical CE (:86-103), 1 MSE average (:105-122), 2 identity (:124-139)
REF_API int ref_dense_loss_backward(void *stream, int kind, float *g, float const *p, float const *y, long long n,
                                    float scale) {
  if (kind == 0)
    ref_loss::categorical_crossentropy_loss_backward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(g, p, y, n);
  else if (kind == 1)
    ref_loss::mean_squared_error_avg_loss_backward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(g, p, y, n);
  else
    ref_loss::identity_loss_backward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(g, p, n);
  ref_helper::scale_kernel<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(g, n, 0, scale);
  return done();
}

// ------------------------------------------------------------------ optimizers
// optimizer_kernel.cu:43-70 (SGD, parameter-server form) and :147-178 (Adam): replica gradients
// are first summed into replica 0, then one update kernel.
REF_API int ref_sgd_update(void *stream, float lr, float momentum, int nesterov, float wd, float *grad, size_t n,
                           int reps, float *W, float *V) {
  for (int i = 1; i < reps; i++)
    ref_helper::apply_add_with_scale<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(grad, grad + i * n, n, 1.0f);
  ref_opt::sgd_update<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, lr, wd, momentum, nesterov != 0, grad, V, W);
  return done();
}
REF_API int ref_adam_update(void *stream, float alpha_t, float b1, float b2, float wd, float eps, float *grad, size_t n,
                            int reps, float *M, float *V, float *W) {
  for (int i = 1; i < reps; i++)
    ref_opt::add_kernel<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>((int)n, 1.0f, grad + i * n, grad);
  ref_opt::adam_update<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>((int)n, alpha_t, b1, b2, wd, eps, grad, M, V, W);
  return done();
}

// ------------------------------------------------------------------ Gather (gather_kernels.cu:80-175)
REF_API int ref_gather_forward(void *stream, float const *in, void const *index, float *out, long long n,
                               long long stride, long long in_dim, long long out_dim, int idx64) {
  if (idx64)
    ref_gather::gather_forward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(in, (int64_t const *)index, out, n, stride, in_dim, out_dim);
  else
    ref_gather::gather_forward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(in, (int32_t const *)index, out, n, stride, in_dim, out_dim);
  return done();
}
REF_API int ref_gather_backward(void *stream, float const *dout, void const *index, float *din, long long n,
                                long long stride, long long in_dim, long long out_dim, int idx64) {
  if (idx64)
    ref_gather::gather_backward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(dout, (int64_t const *)index, din, n, stride, in_dim, out_dim);
  else
    ref_gather::gather_backward<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(dout, (int32_t const *)index, din, n, stride, in_dim, out_dim);
  return done();
}

// ------------------------------------------------------------------ Transpose
// transpose_kernels.cu:45-60, the kernel body only: strides and permutation are passed in LEGION
// order (dim 0 innermost) as the kernel reads them. The reference's host code (:62-120) derives
// the strides with a `+ 1` slip, so the caller passes the INTENDED strides; `out`
// must be zero-filled (the kernel computes out += out * beta + in).
REF_API int ref_transpose_kernel(void *stream, float const *in, float *out, long long volume, int num_dim,
                                 int const *in_strides, int const *out_strides, int const *perm, float beta) {
  ref_transpose::TransposeStrides info;
  info.num_dim = num_dim;
  for (int i = 0; i < num_dim; ++i) {
    info.in_strides[i] = in_strides[i];
    info.out_strides[i] = out_strides[i];
    info.perm[i] = perm[i];
  }
  ref_transpose::transpose_simple_kernel<<<GET_BLOCKS(volume), CUDA_NUM_THREADS, 0, S(stream)>>>(volume, in, out, info, beta);
  return done();
}

// ------------------------------------------------------------------ Cast (cast_kernels.cu:39-59)
// kind: 0 float->half, 1 half->float, 2 int32->float, 3 float->int32, 4 float->double, 5 int64->float
REF_API int ref_cast_forward(void *stream, int kind, void const *in, void *out, size_t n) {
  cudaStream_t s = S(stream);
  int g = GET_BLOCKS(n);
  switch (kind) {
    case 0: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((float const *)in, (half *)out, n); break;
    case 1: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((half const *)in, (float *)out, n); break;
    case 2: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((int32_t const *)in, (float *)out, n); break;
    case 3: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((float const *)in, (int32_t *)out, n); break;
    case 4: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((float const *)in, (double *)out, n); break;
    default: ref_cast::cast_forward<<<g, CUDA_NUM_THREADS, 0, s>>>((int64_t const *)in, (float *)out, n); break;
  }
  return done();
}

// ------------------------------------------------------------------ helper kernels (cuda_helper.cu)
// which: 0 reluBackward (:58-63, launched at :65-82), 1 sigmoid_backward_function (:84-90, :92-109)
REF_API int ref_activation_backward(void *stream, int which, float *grad, float const *output, size_t n) {
  if (which == 0)
    ref_helper::reluBackward<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(grad, output, n);
  else
    ref_helper::sigmoid_backward_function<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(grad, output, n);
  return done();
}
// gelu_forward_kernel (:111-120) as Linear launches it (linear_kernels.cu:180-186), in place
REF_API int ref_gelu_forward_inplace(void *stream, float *x, size_t n) {
  constexpr float B = 0.7978845608028654f, C = 0.035677408136300125f;
  ref_helper::gelu_forward_kernel<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, B, C, x);
  return done();
}
REF_API int ref_copy_with_stride(void *stream, float *out, float const *in, int num_blocks, int out_blk, int in_blk, int add) {
  int n = num_blocks * (out_blk < in_blk ? out_blk : in_blk);
  if (add)
    ref_helper::add_with_stride<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(out, in, num_blocks, out_blk, in_blk);
  else
    ref_helper::copy_with_stride<<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(out, in, num_blocks, out_blk, in_blk);
  return done();
}

// ------------------------------------------------------------------ ElementUnary custom ops
// element_unary_kernels.cu:247-337: ops cuDNN does not serve. op codes are this file's
// OperatorType order. scalar ops use the *_scalar_* kernels.
static bool is_scalar_op(int op) { return op >= (int)OperatorType::SCALAR_MULTIPLY; }
REF_API int ref_element_unary_forward(void *stream, int op, float scalar, float const *x, float *y, long long n) {
  if (is_scalar_op(op))
    ref_unary::elewise_scalar_unary_forward_kernel<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, scalar, (OperatorType)op, x, y);
  else
    ref_unary::elewise_unary_forward_kernel<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, (OperatorType)op, x, y);
  return done();
}
REF_API int ref_element_unary_backward(void *stream, int op, float scalar, float const *y, float const *dy,
                                       float const *x, float *dx, long long n) {
  if (is_scalar_op(op))
    ref_unary::elewise_scalar_unary_backward_kernel<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, scalar, (OperatorType)op, y, dy, x, dx);
  else
    ref_unary::elewise_unary_backward_kernel<float><<<GET_BLOCKS(n), CUDA_NUM_THREADS, 0, S(stream)>>>(n, (OperatorType)op, y, dy, x, dx);
  return done();
}

// ------------------------------------------------------------------ metrics (metrics_functions.cu:139-183)
// flags: bit 0 accuracy, 1 categorical CE, 2 sparse categorical CE, 3 MSE, 4 RMSE, 5 MAE.
// perf: device struct {int all, correct; float cce, sparse_cce, mse, rmse, mae}.
static Metrics metrics_from(int f) {
  return Metrics{(f & 1) != 0, (f & 2) != 0, (f & 4) != 0, (f & 8) != 0, (f & 16) != 0, (f & 32) != 0};
}
REF_API int ref_metrics_sparse(void *stream, int flags, float const *logits, int const *labels, int samples, int classes, void *perf) {
  ref_metrics::update_metrics_sparse_label_kernel<<<GET_BLOCKS(samples), CUDA_NUM_THREADS, 0, S(stream)>>>(
      logits, labels, (PerfMetrics *)perf, metrics_from(flags), samples, classes);
  return done();
}
REF_API int ref_metrics_label(void *stream, int flags, float const *logits, float const *labels, int samples, int classes, void *perf) {
  ref_metrics::update_metrics_label_kernel<<<GET_BLOCKS(samples), 256, 0, S(stream)>>>(
      logits, labels, (PerfMetrics *)perf, metrics_from(flags), samples, classes);
  return done();
}

// ------------------------------------------------------------------ TopK (topk_kernels.cu:372-437)
REF_API int ref_topk_forward(void *stream, float const *in, float *out, int *indices, size_t batch, int length, int k, int sorted) {
  int num_shards = 0;
  {
    constexpr auto shared_memory_size = 48 << 10;
    auto const heap_size = k * sizeof(ref_topk::Entry<float>);
    num_shards = shared_memory_size / heap_size - 1;
    if (num_shards <= 0) return -2;
    if (num_shards > CUDA_NUM_THREADS) num_shards = CUDA_NUM_THREADS;
  }
  size_t shared_memory_size = (num_shards + 1) * k * sizeof(ref_topk::Entry<float>);
  if (num_shards < k) return -2;
  num_shards = k;
  ref_topk::topk_forward_kernel<<<batch, num_shards, 0, S(stream)>>>(in, shared_memory_size, length, k, sorted != 0, out, indices);
  return done();
}
REF_API int ref_topk_backward(void *stream, float const *dout, int const *indices, float *din, size_t batch, int length, int k) {
  ref_topk::topk_backward_kernel<<<GET_BLOCKS(batch * k), CUDA_NUM_THREADS, 0, S(stream)>>>(dout, indices, din, batch, length, k);
  return done();
}

// ------------------------------------------------------------------ Replicate bwd / Reduction fwd
// replicate_kernels.cu:50-64, reduction_kernels.cu:37-52 (grid sized by elements x replicas)
REF_API int ref_replicate_backward(void *stream, float *in_grad, float const *out_grad, size_t n, size_t reps) {
  ref_replicate::replicate_backward_kernel<float><<<GET_BLOCKS(n * reps), CUDA_NUM_THREADS, 0, S(stream)>>>(in_grad, out_grad, n, reps);
  return done();
}
REF_API int ref_reduction_forward(void *stream, float const *in, float *out, size_t n, size_t reps) {
  ref_reduction::reduction_forward_kernel<float><<<GET_BLOCKS(n * reps), CUDA_NUM_THREADS, 0, S(stream)>>>(in, out, n, reps);
  return done();
}

// ------------------------------------------------------------------ cuBLAS calls of Linear / BatchMatmul
static cublasHandle_t blas() {
  static cublasHandle_t h = nullptr;
  if (!h) cublasCreate(&h);
  return h;
}
// linear_kernels.cu:127-149: Y[batch][out] = X[batch][in] * W[out][in]^T as cublasGemmEx(OP_T, OP_N,
// out, batch, in), alpha 1, beta 0, CUBLAS_GEMM_DEFAULT_TENSOR_OP. dtype: 0 fp32 (compute 32F —
// the reference asks for COMPUTE_16F, which cuBLAS rejects for fp32 data), 1 half with
// CUBLAS_COMPUTE_16F as the reference, 2 bf16 with COMPUTE_32F (no bf16 in the reference).
REF_API int ref_linear_forward_gemm(void *stream, void const *x, void const *w, void *y, int in_dim, int out_dim, int batch, int dtype) {
  cublasSetStream(blas(), S(stream));
  cudaDataType_t t = dtype == 0 ? CUDA_R_32F : dtype == 1 ? CUDA_R_16F : CUDA_R_16BF;
  cublasStatus_t st;
  if (dtype == 1) {
    __half alpha = __float2half(1.f), beta = __float2half(0.f);
    st = cublasGemmEx(blas(), CUBLAS_OP_T, CUBLAS_OP_N, out_dim, batch, in_dim, &alpha, w, t, in_dim, x, t, in_dim, &beta,
                      y, t, out_dim, CUBLAS_COMPUTE_16F, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
  } else {
    float alpha = 1.f, beta = 0.f;
    st = cublasGemmEx(blas(), CUBLAS_OP_T, CUBLAS_OP_N, out_dim, batch, in_dim, &alpha, w, t, in_dim, x, t, in_dim, &beta,
                      y, t, out_dim, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
  }
  return st == CUBLAS_STATUS_SUCCESS ? 0 : -(int)st - 100;
}
// batch_matmul_kernels.cu:62-83: O[b] (n x m) = A[b] (n x k) * B[b] (k x m), row-major, as
// cublasSgemmStridedBatched(OP_N, OP_N, m, n, k, B, ldb = m, A, lda = k, O, ldo = m)
REF_API int ref_batch_matmul_forward(void *stream, float *o, float const *a, float const *b, int m, int n, int k, int batch) {
  cublasSetStream(blas(), S(stream));
  float alpha = 1.f, beta = 0.f;
  cublasStatus_t st = cublasSgemmStridedBatched(blas(), CUBLAS_OP_N, CUBLAS_OP_N, m, n, k, &alpha, b, m, (long long)k * m,
                                                a, k, (long long)n * k, &beta, o, m, (long long)n * m, batch);
  return st == CUBLAS_STATUS_SUCCESS ? 0 : -(int)st - 100;
}

// linear_kernels.cu:229-327 — the three products of Linear's backward, all with alpha = beta = 1 (the
// gradients ACCUMULATE), column-major views of the row-major tensors:
//   dW[out][in] += dY[batch][out]^T X[batch][in]     cublasGemmEx(OP_N, OP_T, in, out, batch, X, in, dY, out, dW, in)
//   db[out]     += 1^T dY                            cublasGemmEx(OP_N, OP_T, 1, out, batch, ones, 1, dY, out, db, 1)
//   dX[batch][in] += dY W[out][in]                   cublasGemmEx(OP_N, OP_N, in, batch, out, W, in, dY, out, dX, in)
// dtype as in ref_linear_forward_gemm. `ones` has `batch` elements OF THE DATA TYPE: the reference
// passes its fp32 one_ptr with CUDA_R_32F beside 16-bit dY, a combination cublasGemmEx rejects
// (CUBLAS_STATUS_NOT_SUPPORTED) — the ledger's deviation for 16-bit bias gradients.
REF_API int ref_linear_backward_gemm(void *stream, void const *x, void *dx, void const *dy, void const *w, void *dw,
                                     void *db, void const *ones, int in_dim, int out_dim, int batch, int dtype) {
  cublasSetStream(blas(), S(stream));
  cudaDataType_t t = dtype == 0 ? CUDA_R_32F : dtype == 1 ? CUDA_R_16F : CUDA_R_16BF;
  cublasComputeType_t ct = dtype == 1 ? CUBLAS_COMPUTE_16F : CUBLAS_COMPUTE_32F;
  __half const h1 = __float2half(1.f);
  float const f1 = 1.f;
  void const *one = dtype == 1 ? static_cast<void const *>(&h1) : static_cast<void const *>(&f1);
  cublasStatus_t st = cublasGemmEx(blas(), CUBLAS_OP_N, CUBLAS_OP_T, in_dim, out_dim, batch, one, x, t, in_dim, dy, t,
                                   out_dim, one, dw, t, in_dim, ct, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
  if (st != CUBLAS_STATUS_SUCCESS) return -(int)st - 100;
  if (db != nullptr) {
    st = cublasGemmEx(blas(), CUBLAS_OP_N, CUBLAS_OP_T, 1, out_dim, batch, one, ones, t, 1, dy, t, out_dim, one, db, t,
                      1, ct, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
    if (st != CUBLAS_STATUS_SUCCESS) return -(int)st - 200;
  }
  if (dx != nullptr) {
    st = cublasGemmEx(blas(), CUBLAS_OP_N, CUBLAS_OP_N, in_dim, batch, out_dim, one, w, t, in_dim, dy, t, out_dim, one,
                      dx, t, in_dim, ct, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
    if (st != CUBLAS_STATUS_SUCCESS) return -(int)st - 300;
  }
  return 0;
}
// batch_matmul_kernels.cu:85-139: dA[b] (n x k) += dO[b] (n x m) B[b]^T (m x k), dB[b] (k x m) += A[b]^T dO[b],
// both cublasSgemmStridedBatched with alpha = beta = 1
REF_API int ref_batch_matmul_backward(void *stream, float const *o_grad, float const *a, float *a_grad, float const *b,
                                      float *b_grad, int m, int n, int k, int batch) {
  cublasSetStream(blas(), S(stream));
  float alpha = 1.f;
  long long const a_stride = (long long)n * k, b_stride = (long long)m * k, o_stride = (long long)n * m;
  cublasStatus_t st = cublasSgemmStridedBatched(blas(), CUBLAS_OP_T, CUBLAS_OP_N, k, n, m, &alpha, b, m, b_stride, o_grad,
                                                m, o_stride, &alpha, a_grad, k, a_stride, batch);
  if (st != CUBLAS_STATUS_SUCCESS) return -(int)st - 100;
  st = cublasSgemmStridedBatched(blas(), CUBLAS_OP_N, CUBLAS_OP_T, m, k, n, &alpha, o_grad, m, o_stride, a, k, a_stride,
                                 &alpha, b_grad, m, b_stride, batch);
  return st == CUBLAS_STATUS_SUCCESS ? 0 : -(int)st - 200;
}
