// C++ veneer, part 3: FlexFlow::Kernels::<Op>::{init,forward,backward}_kernel and the loss /
// optimizer / metrics functions, with the reference's names and argument order, as inline
// forwards to the C ABI (include/ffk*.h). Per-op state structs keep the reference's names;
// instead of cuDNN descriptors they carry the plain numbers the sm_100a kernels need.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstring>
#include <stdexcept>
#include <variant>
#include <vector>
#include "runtime.h"

namespace FlexFlow {

inline int ffk_dt(DataType dt) { return static_cast<int>(dt); }
inline int ffk_op(OperatorType op) { return static_cast<int>(op); }

// ------------------------------------------------------------------ per-op states
struct LinearPerDeviceState {
  PerDeviceFFHandle handle;
  float const *one_ptr;  // kept for signature compatibility; unused (the bias is fused)
  std::optional<Activation> activation;
  std::optional<RegularizerAttrs> regularizer;
  bool use_bias;
  DataType input_type, weight_type, output_type;
};
struct MHAPerDeviceState {
  PerDeviceFFHandle handle;
  size_t weightSize;
  size_t reserveSpaceSize;
  ffk_mha_config_t config;
  void *reserveSpace;
  void *workSpace;  // backward scratch (cuDNN used handle.workSpace; ours is sized per config)
  size_t workSpaceSize;
  Allocator allocator;
};
struct ElementUnaryPerDeviceState {
  ArrayShape input_shape, output_shape;
  OperatorType op_type;
};
struct ElementBinaryPerDeviceState {
  PerDeviceFFHandle handle;
  OperatorType op_type;
  ArrayShape lhs_shape, rhs_shape, output_shape;
};
struct SoftmaxPerDeviceState {
  PerDeviceFFHandle handle;
  int dim;
  int n, c, h, w;
};
struct LayerNormPerDeviceState {
  PerDeviceFFHandle handle;
  bool elementwise_affine;
  int64_t effective_batch_size, effective_num_elements;
  float eps;
  float *mean, *rstd, *ds, *db, *scale, *bias;
  DataType data_type;
};
// cuDNN descriptor handles of the reference's state structs (include/kernels/device.h typedefs):
// kept as opaque pointers so that the member layout and the cleanup_kernel signatures stay the
// reference's; nothing here creates or reads a cuDNN descriptor, the slots are null.
using ffTensorDescriptor_t = struct ffk_opaque_tensor_descriptor *;
using ffActivationDescriptor_t = struct ffk_opaque_activation_descriptor *;
using ffDropoutDescriptor_t = struct ffk_opaque_dropout_descriptor *;
using ffBatchNormMode_t = int;

// batch_norm_kernels.h:11-27, member for member
struct BatchNormPerDeviceState {
  PerDeviceFFHandle handle;
  ffTensorDescriptor_t inputTensor, outputTensor, biasTensor;
  ffActivationDescriptor_t actiDesc;
  ffBatchNormMode_t mode;
  float *runningMean, *runningVar, *saveMean, *saveVar;  // saveVar holds 1/sqrt(var+eps)
  int output_n, output_c, output_h, output_w;
  bool relu;
};
// dropout_kernels.h:12-22, member for member; the generator state of this library follows them
struct DropoutPerDeviceState {
  PerDeviceFFHandle handle;
  ffTensorDescriptor_t inputTensor, outputTensor;
  ffDropoutDescriptor_t dropoutDesc;
  void *reserveSpace;                 // 1 bit per element keep-mask
  void *dropoutStates;                // unused: the Philox stream is keyed by (seed, offset)
  size_t reserveSpaceSize;
  size_t dropoutStateSize;
  float rate;
  unsigned long long seed;
  mutable unsigned long long offset;  // advances once per forward call (fresh mask each step)
  size_t num_elements;
};
static_assert(offsetof(BatchNormPerDeviceState, runningMean) ==
                  sizeof(PerDeviceFFHandle) + 4 * sizeof(void *) + sizeof(void *),
              "BatchNormPerDeviceState: descriptors and mode sit between the handle and runningMean, as in the reference");
static_assert(offsetof(DropoutPerDeviceState, reserveSpace) == sizeof(PerDeviceFFHandle) + 3 * sizeof(void *),
              "DropoutPerDeviceState: three descriptor slots between the handle and reserveSpace, as in the reference");
struct ReducePerDeviceState {
  PerDeviceFFHandle handle;
  OperatorType op_type;
  size_t reduction_size;
  ArrayShape input_shape, output_shape;
};
struct Pool2DPerDeviceState {
  PerDeviceFFHandle handle;
  int n, c, in_h, in_w, out_h, out_w, kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w;
  PoolOp pool_type;
  bool relu;
};
struct Conv2DPerDeviceState {
  PerDeviceFFHandle handle;
  ffk_conv2d_config_t config;
};
struct GatherPerDeviceState {
  PerDeviceFFHandle handle;
  legion_dim_t legion_dim;
};
struct TransposePerDeviceState {
  int num_dim;
  std::vector<ff_dim_t> perm;  // output FF dim i takes input FF dim perm[i]
};
struct TopKPerDeviceState { bool sorted; };
struct ReshapePerDeviceState { DataType data_type; };
struct RepartitionPerDeviceState {
  PerDeviceFFHandle handle;
  DataType data_type;
};

namespace Kernels {

// ------------------------------------------------------------------ linear_kernels.h
namespace Linear {
inline bool use_activation(Activation a) {
  return a == Activation::RELU || a == Activation::SIGMOID || a == Activation::TANH;
}
inline LinearPerDeviceState init_kernel(PerDeviceFFHandle handle, float *one_ptr,
                                        std::optional<Activation> activation,
                                        std::optional<RegularizerAttrs> regularizer, bool use_bias,
                                        DataType input_type, DataType weight_type,
                                        DataType output_type, int /*batch_size*/, int /*channel*/) {
  if (input_type != weight_type || input_type != output_type)
    throw mk_runtime_error("Linear: input, weight and output must share one data type");
  return LinearPerDeviceState{handle, one_ptr, activation, regularizer, use_bias,
                              input_type, weight_type, output_type};
}
inline void forward_kernel(ffStream_t stream, LinearPerDeviceState const &m, void const *input_ptr,
                           void *output_ptr, void const *filter_ptr, void const *bias_ptr,
                           int in_dim, int out_dim, int batch_size) {
  FFK_CHECK(ffk_linear_forward(stream, input_ptr, output_ptr, filter_ptr, bias_ptr, in_dim, out_dim,
                               batch_size, ffk_dt(m.input_type), act_code(m.activation),
                               FFK_GEMM_AUTO));
}
inline void backward_kernel(ffStream_t stream, LinearPerDeviceState const &m, void const *input_ptr,
                            void *input_grad_ptr, void const *output_ptr, void *output_grad_ptr,
                            void const *kernel_ptr, void *kernel_grad_ptr, void *bias_ptr,
                            int in_dim, int out_dim, int batch_size) {
  float lambda = 0.f;
  if (m.regularizer.has_value()) {
    if (!m.regularizer->has<L2RegularizerAttrs>()) FatalError("Only L2 regularization is supported");
    lambda = m.regularizer->get<L2RegularizerAttrs>().lambda;
  }
  FFK_CHECK(ffk_linear_backward(stream, input_ptr, input_grad_ptr, output_ptr, output_grad_ptr,
                                kernel_ptr, kernel_grad_ptr, bias_ptr, in_dim, out_dim, batch_size,
                                ffk_dt(m.input_type), act_code(m.activation), lambda, nullptr,
                                FFK_GEMM_AUTO, m.handle.workSpace, m.handle.workSpaceSize));
}
}  // namespace Linear

// ------------------------------------------------------------------ batch_matmul_kernels.h
namespace BatchMatmul {
// argument order of the reference DEFINITION (batch_matmul_kernels.cu:23-37), which its
// callers use (lib/local-execution/src/ops/batch_matmul.cc:97-99)
inline void forward_kernel(ffStream_t stream, PerDeviceFFHandle const &, float *output_ptr,
                           float const *a_input_ptr, float const *b_input_ptr, int m, int n, int k,
                           int batch, int a_seq_length_dim, int b_seq_length_dim, int seq_length) {
  FFK_CHECK(ffk_batch_matmul_forward(stream, output_ptr, a_input_ptr, b_input_ptr, m, n, k, batch,
                                     a_seq_length_dim, b_seq_length_dim, seq_length, FFK_FLOAT,
                                     FFK_GEMM_AUTO));
}
inline void backward_kernel(ffStream_t stream, PerDeviceFFHandle const &, float const *o_ptr,
                            float const *o_grad_ptr, float const *a_ptr, float *a_grad_ptr,
                            float const *b_ptr, float *b_grad_ptr, int m, int n, int k, int batch) {
  FFK_CHECK(ffk_batch_matmul_backward(stream, o_ptr, o_grad_ptr, a_ptr, a_grad_ptr, b_ptr,
                                      b_grad_ptr, m, n, k, batch, FFK_FLOAT, FFK_GEMM_AUTO));
}
}  // namespace BatchMatmul

// ------------------------------------------------------------------ attention_kernels.h
namespace MultiHeadAttention {
inline MHAPerDeviceState init_kernel(PerDeviceFFHandle const &handle, Allocator &allocator,
                                     int num_samples, int num_heads, int qSize, int kSize, int vSize,
                                     int qProjSize, int kProjSize, int vProjSize, int oProjSize,
                                     int qoSeqLength, int kvSeqLength, bool add_bias_kv) {
  if (add_bias_kv) FatalError("MultiHeadAttention: add_bias_kv is not supported");
  ffk_mha_config_t c{num_samples, num_heads, qSize, kSize, vSize, qProjSize, kProjSize,
                     vProjSize, oProjSize, qoSeqLength, kvSeqLength};
  size_t wsz = ffk_mha_weight_bytes(&c, FFK_FLOAT);
  if (wsz == 0) FatalError(std::string("MultiHeadAttention: ") + ffk_last_error());
  size_t rsz = ffk_mha_reserve_bytes(&c, FFK_FLOAT), ssz = ffk_mha_workspace_bytes(&c, FFK_FLOAT);
  void *reserve = allocator.allocate(rsz ? rsz : 16);
  void *work = allocator.allocate(ssz ? ssz : 16);
  return MHAPerDeviceState{handle, wsz, rsz, c, reserve, work, ssz, allocator};
}
inline void forward_kernel(ffStream_t stream, MHAPerDeviceState const &s, float const *query_ptr,
                           float const *key_ptr, float const *value_ptr, float const *weight_ptr,
                           float *output_ptr) {
  FFK_CHECK(ffk_mha_forward(stream, &s.config, query_ptr, key_ptr, value_ptr, weight_ptr,
                            output_ptr, s.reserveSpace, FFK_FLOAT, FFK_GEMM_AUTO));
}
inline void backward_kernel(ffStream_t stream, MHAPerDeviceState const &s, float const *query_ptr,
                            float *query_grad_ptr, float const *key_ptr, float *key_grad_ptr,
                            float const *value_ptr, float *value_grad_ptr, float const *weight_ptr,
                            float *weight_grad_ptr, float const *output_grad_ptr) {
  FFK_CHECK(ffk_mha_backward(stream, &s.config, query_ptr, query_grad_ptr, key_ptr, key_grad_ptr,
                             value_ptr, value_grad_ptr, weight_ptr, weight_grad_ptr,
                             output_grad_ptr, s.reserveSpace, s.workSpace, s.workSpaceSize,
                             FFK_FLOAT, FFK_GEMM_AUTO));
}
inline void cleanup_kernel(Allocator &allocator, MHAPerDeviceState const &s) {
  allocator.deallocate(s.reserveSpace);
  allocator.deallocate(s.workSpace);
}
}  // namespace MultiHeadAttention

// ------------------------------------------------------------------ element_unary_kernels.h
namespace ElementUnary {
inline ElementUnaryPerDeviceState init_kernel(ArrayShape const &input_shape,
                                              ArrayShape const &output_shape,
                                              ElementUnaryAttrs const &attrs) {
  return ElementUnaryPerDeviceState{input_shape, output_shape, get_op_type(attrs)};
}
inline void forward_kernel(ffStream_t stream, ElementUnaryPerDeviceState const &,
                           ElementUnaryAttrs const &attrs, PerDeviceFFHandle const &,
                           GenericTensorAccessorR const &input, GenericTensorAccessorW const &output) {
  FFK_CHECK(ffk_element_unary_forward(stream, ffk_op(attrs.op_type), attrs.scalar.value_or(0.f),
                                      input.ptr, output.ptr, input.shape.num_elements(),
                                      ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, ElementUnaryPerDeviceState const &,
                            ElementUnaryAttrs const &attrs, PerDeviceFFHandle const &,
                            GenericTensorAccessorR const &input, GenericTensorAccessorW const &input_grad,
                            GenericTensorAccessorR const &output, GenericTensorAccessorR const &output_grad) {
  FFK_CHECK(ffk_element_unary_backward(stream, ffk_op(attrs.op_type), attrs.scalar.value_or(0.f),
                                       input.ptr, input_grad.ptr, output.ptr, output_grad.ptr,
                                       input.shape.num_elements(), ffk_dt(input.data_type)));
}
}  // namespace ElementUnary

// ------------------------------------------------------------------ element_binary_kernels.h
namespace ElementBinary {
inline ElementBinaryPerDeviceState init_kernel(PerDeviceFFHandle handle, OperatorType op_type,
                                               bool /*should_broadcast_lhs*/, bool /*should_broadcast_rhs*/,
                                               ArrayShape lhs_shape, ArrayShape rhs_shape,
                                               ArrayShape output_shape) {
  return ElementBinaryPerDeviceState{handle, op_type, lhs_shape, rhs_shape, output_shape};
}
namespace detail {
// shapes of different rank are right-aligned in FF order, as NCHW descriptors would be
inline void padded(ElementBinaryPerDeviceState const &m, std::vector<long long> &l,
                   std::vector<long long> &r, std::vector<long long> &o) {
  o = m.output_shape.ff_ordered();
  auto pad = [&](ArrayShape const &s) {
    std::vector<long long> v = s.ff_ordered();
    v.insert(v.begin(), o.size() - v.size(), 1);
    return v;
  };
  l = pad(m.lhs_shape);
  r = pad(m.rhs_shape);
}
}  // namespace detail
inline void forward_kernel(ffStream_t stream, ElementBinaryPerDeviceState const &m,
                           float const *lhs_ptr, float const *rhs_ptr, float *out_ptr,
                           OperatorType op_type, bool /*broadcast_inputLHS*/, PerDeviceFFHandle) {
  std::vector<long long> l, r, o;
  detail::padded(m, l, r, o);
  FFK_CHECK(ffk_element_binary_forward(stream, ffk_op(op_type), lhs_ptr, rhs_ptr, out_ptr,
                                       (int)o.size(), l.data(), r.data(), o.data(), FFK_FLOAT));
}
inline void backward_kernel(ffStream_t stream, ElementBinaryPerDeviceState const &m,
                            float const *out_grad_ptr, float const *lhs_ptr, float const *rhs_ptr,
                            float *lhs_grad_ptr, float *rhs_grad_ptr, OperatorType op_type,
                            bool, bool, PerDeviceFFHandle) {
  std::vector<long long> l, r, o;
  detail::padded(m, l, r, o);
  FFK_CHECK(ffk_element_binary_backward(stream, ffk_op(op_type), out_grad_ptr, lhs_ptr, rhs_ptr,
                                        lhs_grad_ptr, rhs_grad_ptr, (int)o.size(), l.data(),
                                        r.data(), o.data(), FFK_FLOAT));
}
}  // namespace ElementBinary

// ------------------------------------------------------------------ softmax_kernels.h
namespace Softmax {
inline SoftmaxPerDeviceState init_kernel(PerDeviceFFHandle const &handle, int dim, int input_n,
                                         int input_c, int input_h, int input_w) {
  return SoftmaxPerDeviceState{handle, dim, input_n, input_c, input_h, input_w};
}
inline void forward_kernel(ffStream_t stream, SoftmaxPerDeviceState const &m, float const *input_ptr,
                           float *output_ptr) {
  FFK_CHECK(ffk_softmax_forward(stream, input_ptr, output_ptr, m.n, m.c, (long long)m.h * m.w, FFK_FLOAT));
}
inline void backward_kernel(ffStream_t stream, float *input_grad_ptr, float const *output_grad_ptr,
                            size_t num_elements) {
  FFK_CHECK(ffk_softmax_backward(stream, input_grad_ptr, output_grad_ptr, nullptr,
                                 (long long)num_elements, 1, 1, FFK_FLOAT, 0));
}
}  // namespace Softmax

// ------------------------------------------------------------------ layer_norm_kernels.h
namespace LayerNorm {
inline LayerNormPerDeviceState init_kernel(PerDeviceFFHandle const &handle, Allocator &allocator,
                                           bool elementwise_affine, int64_t effective_batch_size,
                                           int64_t effective_num_elements, float eps) {
  auto f = [&]() { return static_cast<float *>(allocator.allocate(sizeof(float) * effective_batch_size)); };
  float *mean = f(), *rstd = f(), *ds = f(), *db = f(), *scale = f(), *bias = f();
  return LayerNormPerDeviceState{handle, elementwise_affine, effective_batch_size,
                                 effective_num_elements, eps, mean, rstd, ds, db, scale, bias,
                                 DataType::FLOAT};
}
inline void forward_kernel(ffStream_t stream, LayerNormPerDeviceState const &m,
                           GenericTensorAccessorR const &input, GenericTensorAccessorW const &output,
                           GenericTensorAccessorW const &gamma, GenericTensorAccessorW const &beta) {
  FFK_CHECK(ffk_layer_norm_forward(stream, input.ptr, output.ptr, gamma.ptr, beta.ptr, m.mean, m.rstd,
                                   m.effective_batch_size, m.effective_num_elements, m.eps,
                                   ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, LayerNormPerDeviceState const &m,
                            GenericTensorAccessorR const &output_grad, GenericTensorAccessorR const &input,
                            GenericTensorAccessorW const &input_grad, GenericTensorAccessorR const &gamma,
                            GenericTensorAccessorW const &gamma_grad, GenericTensorAccessorW const &beta_grad) {
  FFK_CHECK(ffk_layer_norm_backward(stream, output_grad.ptr, input.ptr, input_grad.ptr, gamma.ptr,
                                    gamma_grad.ptr, beta_grad.ptr, m.mean, m.rstd,
                                    m.effective_batch_size, m.effective_num_elements,
                                    ffk_dt(input.data_type), 0, m.handle.workSpace,
                                    m.handle.workSpaceSize));
}
}  // namespace LayerNorm

// ------------------------------------------------------------------ batch_norm_kernels.h
namespace BatchNorm {
inline BatchNormPerDeviceState init_kernel(PerDeviceFFHandle handle, Allocator allocator,
                                           float *runningMean, int output_n, int output_c,
                                           int output_h, int output_w, bool relu) {
  runningMean = static_cast<float *>(allocator.allocate(sizeof(float) * output_c * 4));
  float *runningVar = runningMean + output_c, *saveMean = runningVar + output_c,
        *saveVar = saveMean + output_c;
  cudaStream_t s;
  checkCUDA(get_legion_stream(&s));
  FFK_CHECK(ffk_fill(s, runningMean, size_t(output_c) * 2, 0.0, FFK_FLOAT));
  return BatchNormPerDeviceState{handle, nullptr, nullptr, nullptr, nullptr, 0, runningMean, runningVar, saveMean,
                                 saveVar, output_n, output_c, output_h, output_w, relu};
}
inline void forward_kernel(ffStream_t stream, BatchNormPerDeviceState const &m, float const *input_ptr,
                           float *output_ptr, float const *scale_ptr, float const *bias_ptr) {
  // eps = CUDNN_BN_MIN_EPSILON (0.0 in cuDNN 9), factor 1.0: batch_norm_kernels.cu:35-52
  FFK_CHECK(ffk_batch_norm_forward(stream, input_ptr, output_ptr, scale_ptr, bias_ptr, m.runningMean,
                                   m.runningVar, m.saveMean, m.saveVar, m.output_n, m.output_c,
                                   m.output_h, m.output_w, 0.0f, 1.0f));
}
inline void backward_kernel(ffStream_t stream, BatchNormPerDeviceState const &m, float const *input_ptr,
                            float *output_grad_ptr, float const *output_ptr, float *input_grad_ptr,
                            float const *scale_ptr, float *scale_grad_ptr, float *bias_grad_ptr,
                            size_t /*numElements*/) {
  FFK_CHECK(ffk_batch_norm_backward(stream, input_ptr, output_grad_ptr, output_ptr, input_grad_ptr,
                                    scale_ptr, scale_grad_ptr, bias_grad_ptr, m.saveMean, m.saveVar,
                                    m.output_n, m.output_c, m.output_h, m.output_w, m.relu));
}
inline void cleanup_kernel(Allocator allocator, BatchNormPerDeviceState const &m) {
  allocator.deallocate(m.runningMean);
}
// batch_norm_kernels.h:76-82: the reference's signature (descriptor handles are ignored here)
inline void cleanup_kernel(Allocator allocator, ffTensorDescriptor_t, ffTensorDescriptor_t, ffTensorDescriptor_t,
                           ffActivationDescriptor_t, bool /*relu*/, float *runningMean) {
  allocator.deallocate(runningMean);
}
}  // namespace BatchNorm

// ------------------------------------------------------------------ dropout_kernels.h
namespace Dropout {
inline DropoutPerDeviceState init_kernel(PerDeviceFFHandle handle, float rate, unsigned long long seed,
                                         ArrayShape const &output_domain, Allocator allocator) {
  size_t n = output_domain.num_elements();
  size_t bytes = ffk_dropout_mask_bytes(n);
  void *mask = allocator.allocate(bytes ? bytes : 4);
  return DropoutPerDeviceState{handle, nullptr, nullptr, nullptr, mask, nullptr, bytes, 0, rate, seed, 0ull, n};
}
inline void forward_kernel(ffStream_t stream, DropoutPerDeviceState const &m, float const *input_ptr,
                           float *output_ptr) {
  FFK_CHECK(ffk_dropout_forward(stream, input_ptr, output_ptr, m.reserveSpace, m.num_elements, m.rate,
                                m.seed, m.offset, FFK_FLOAT));
  m.offset += 1;
}
inline void backward_kernel(ffStream_t stream, DropoutPerDeviceState const &m,
                            float const *output_grad_ptr, float *input_grad_ptr) {
  FFK_CHECK(ffk_dropout_backward(stream, output_grad_ptr, input_grad_ptr, m.reserveSpace,
                                 m.num_elements, m.rate, FFK_FLOAT));
}
inline void cleanup_kernel(Allocator allocator, DropoutPerDeviceState const &m) {
  allocator.deallocate(m.reserveSpace);
}
// dropout_kernels.h:53-57: the reference's signature. The reference frees `dropoutStates` here; this
// library keeps no generator state, and the mask (reserveSpace) stays with the allocator.
inline void cleanup_kernel(Allocator allocator, ffTensorDescriptor_t, ffTensorDescriptor_t, ffDropoutDescriptor_t,
                           void *dropoutStates) {
  if (dropoutStates) allocator.deallocate(dropoutStates);
}
}  // namespace Dropout

// ------------------------------------------------------------------ embedding_kernels.h
namespace Embedding {
inline int aggr_code(std::optional<AggregateOp> const &a) {
  return a.has_value() ? static_cast<int>(*a) : FFK_AGGR_NONE;
}
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                           GenericTensorAccessorW const &output, GenericTensorAccessorR const &weight,
                           DataType input_data_type, DataType output_data_type,
                           std::optional<AggregateOp> aggr, int in_dim, int out_dim, int batch_size) {
  FFK_CHECK(ffk_embedding_forward(stream, input.ptr, output.ptr, weight.ptr, ffk_dt(input_data_type),
                                  ffk_dt(output_data_type), aggr_code(aggr), in_dim, out_dim, batch_size));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                            GenericTensorAccessorR const &output, GenericTensorAccessorW const &weight_grad,
                            DataType input_data_type, DataType output_data_type,
                            std::optional<AggregateOp> aggr, int in_dim, int out_dim, int batch_size) {
  FFK_CHECK(ffk_embedding_backward(stream, input.ptr, output.ptr, weight_grad.ptr,
                                   ffk_dt(input_data_type), ffk_dt(output_data_type), aggr_code(aggr),
                                   in_dim, out_dim, batch_size));
}
inline void rand_generate_int64_wrapper(int64_t *ptr, size_t size, int64_t p) {
  cudaStream_t s;
  checkCUDA(get_legion_stream(&s));
  FFK_CHECK(ffk_rand_generate_int(s, ptr, size, p, FFK_INT64));
}
inline void rand_generate_int32_wrapper(int32_t *ptr, size_t size, int32_t p) {
  cudaStream_t s;
  checkCUDA(get_legion_stream(&s));
  FFK_CHECK(ffk_rand_generate_int(s, ptr, size, p, FFK_INT32));
}
}  // namespace Embedding

// ------------------------------------------------------------------ reduce_kernels.h
namespace Reduce {
inline ReducePerDeviceState init_kernel(PerDeviceFFHandle const &handle, OperatorType const &op_type,
                                        size_t const &reduction_size, ArrayShape const &input_shape,
                                        ArrayShape const &output_shape) {
  return ReducePerDeviceState{handle, op_type, reduction_size, input_shape, output_shape};
}
inline void forward_kernel(ffStream_t stream, ReducePerDeviceState const &m, float const *input_ptr,
                           float *output_ptr) {
  auto i = m.input_shape.ff_ordered(), o = m.output_shape.ff_ordered();
  o.insert(o.begin(), i.size() - o.size(), 1);
  FFK_CHECK(ffk_reduce_forward(stream, input_ptr, output_ptr, (int)i.size(), i.data(), o.data(),
                               ffk_op(m.op_type), FFK_FLOAT));
}
inline void backward_kernel(ffStream_t stream, ReducePerDeviceState const &m,
                            float const *output_grad_ptr, float *input_grad_ptr) {
  auto i = m.input_shape.ff_ordered(), o = m.output_shape.ff_ordered();
  o.insert(o.begin(), i.size() - o.size(), 1);
  FFK_CHECK(ffk_reduce_backward(stream, output_grad_ptr, input_grad_ptr, (int)i.size(), i.data(),
                                o.data(), ffk_op(m.op_type), FFK_FLOAT));
}
}  // namespace Reduce

// ------------------------------------------------------------------ pool_2d_kernels.h
namespace Pool2D {
inline Pool2DPerDeviceState init_kernel(PerDeviceFFHandle handle, std::optional<Activation> activation,
                                        int input_w, int input_h, int input_c, int input_n,
                                        int output_w, int output_h, int output_c, int output_n,
                                        int pad_h, int pad_w, int kernel_h, int kernel_w, int stride_h,
                                        int stride_w, PoolOp pool_type) {
  if (output_n != input_n || output_c != input_c ||
      output_h != (input_h + 2 * pad_h - kernel_h) / stride_h + 1 ||
      output_w != (input_w + 2 * pad_w - kernel_w) / stride_w + 1)
    FatalError("Pool2D: output shape does not match the pooling geometry");
  return Pool2DPerDeviceState{handle, input_n, input_c, input_h, input_w, output_h, output_w,
                              kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w, pool_type,
                              activation == Activation::RELU};
}
// pool_2d_kernels.h:49-64 — the in-place form: the state's handle and ReLU flag stay as they are
inline void init_kernel(Pool2DPerDeviceState *m, int input_w, int input_h, int input_c, int input_n, int output_w,
                        int output_h, int output_c, int output_n, int pad_h, int pad_w, int kernel_h, int kernel_w,
                        int stride_h, int stride_w, PoolOp pool_type) {
  *m = init_kernel(m->handle, m->relu ? std::optional<Activation>(Activation::RELU) : std::nullopt, input_w, input_h,
                   input_c, input_n, output_w, output_h, output_c, output_n, pad_h, pad_w, kernel_h, kernel_w,
                   stride_h, stride_w, pool_type);
}
inline void forward_kernel(ffStream_t stream, Pool2DPerDeviceState const &m, void const *input_ptr,
                           void *output_ptr) {
  FFK_CHECK(ffk_pool2d_forward(stream, (float const *)input_ptr, (float *)output_ptr, m.n, m.c, m.in_h,
                               m.in_w, m.out_h, m.out_w, m.kernel_h, m.kernel_w, m.stride_h,
                               m.stride_w, m.pad_h, m.pad_w, static_cast<int>(m.pool_type)));
}
inline void backward_kernel(ffStream_t stream, Pool2DPerDeviceState const &m, void const *input_ptr,
                            void *input_grad_ptr, void const *output_ptr, void const *output_grad_ptr) {
  FFK_CHECK(ffk_pool2d_backward(stream, (float const *)input_ptr, (float *)input_grad_ptr,
                                (float const *)output_ptr, (float const *)output_grad_ptr, m.n, m.c,
                                m.in_h, m.in_w, m.out_h, m.out_w, m.kernel_h, m.kernel_w, m.stride_h,
                                m.stride_w, m.pad_h, m.pad_w, static_cast<int>(m.pool_type)));
}
}  // namespace Pool2D

// ------------------------------------------------------------------ conv_2d_kernels.h
namespace Conv2D {
// input / output accessors are NCHW in FF order, i.e. legion dims (w, h, c, n)
inline Conv2DPerDeviceState init_kernel(PerDeviceFFHandle handle, std::optional<Activation> /*activation*/,
                                        int kernel_h, int kernel_w, int groups, int padding_h,
                                        int padding_w, int stride_h, int stride_w,
                                        GenericTensorAccessorW const &input,
                                        GenericTensorAccessorW const &output, float const * /*filter_ptr*/,
                                        float * /*filter_grad_ptr*/) {
  auto L = [](ArrayShape const &s, int i) { return (int)s[legion_dim_t{i}]; };
  ffk_conv2d_config_t c{L(input.shape, 3), L(input.shape, 2), L(input.shape, 1), L(input.shape, 0),
                        L(output.shape, 2), L(output.shape, 1), L(output.shape, 0),
                        kernel_h, kernel_w, padding_h, padding_w, stride_h, stride_w, groups};
  if (ffk_conv2d_workspace_bytes(&c) > handle.workSpaceSize)
    FatalError("Conv2D: handle.workSpace is smaller than one im2col panel");
  return Conv2DPerDeviceState{handle, c};
}
inline void forward_kernel(ffStream_t stream, Conv2DPerDeviceState const &m, float const *input_ptr,
                           float *output_ptr, float const *filter_ptr, float const *bias_ptr,
                           std::optional<Activation> activation) {
  FFK_CHECK(ffk_conv2d_forward(stream, &m.config, input_ptr, output_ptr, filter_ptr, bias_ptr,
                               activation.has_value(), m.handle.workSpace, m.handle.workSpaceSize));
}
inline void backward_kernel(ffStream_t stream, Conv2DPerDeviceState const &m, float const *input_ptr,
                            float *input_grad_ptr, float const *output_ptr, float *output_grad_ptr,
                            float const *filter_ptr, float *filter_grad_ptr, float *bias_grad_ptr,
                            std::optional<Activation> activation) {
  FFK_CHECK(ffk_conv2d_backward(stream, &m.config, input_ptr, input_grad_ptr, output_ptr, output_grad_ptr,
                                filter_ptr, filter_grad_ptr, bias_grad_ptr, activation.has_value(),
                                m.handle.workSpace, m.handle.workSpaceSize));
}
}  // namespace Conv2D

// ------------------------------------------------------------------ cast / flat / reshape
namespace Cast {
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                           GenericTensorAccessorW const &output, DataType input_type, DataType output_type) {
  FFK_CHECK(ffk_cast_forward(stream, input.ptr, output.ptr, input.shape.num_elements(),
                             ffk_dt(input_type), ffk_dt(output_type)));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                            GenericTensorAccessorW const &output, DataType input_type, DataType output_type) {
  FFK_CHECK(ffk_cast_backward(stream, input.ptr, output.ptr, input.shape.num_elements(),
                              ffk_dt(input_type), ffk_dt(output_type)));
}
}  // namespace Cast
namespace Flat {
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR input, float *output_ptr) {
  FFK_CHECK(ffk_copy(stream, output_ptr, input.ptr, input.shape.num_elements(), ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorR input, float *input_grad_ptr,
                            float const *output_grad_ptr) {
  FFK_CHECK(ffk_accumulate(stream, input_grad_ptr, output_grad_ptr, input.shape.num_elements(), FFK_FLOAT));
}
}  // namespace Flat
namespace Reshape {
inline ReshapePerDeviceState init_kernel(DataType data_type) { return ReshapePerDeviceState{data_type}; }
inline void forward_kernel(ffStream_t stream, ReshapePerDeviceState const &m,
                           GenericTensorAccessorR const &input, GenericTensorAccessorW const &output) {
  FFK_CHECK(ffk_copy(stream, output.ptr, input.ptr, input.shape.num_elements(), ffk_dt(m.data_type)));
}
inline void backward_kernel(ffStream_t stream, ReshapePerDeviceState const &m,
                            GenericTensorAccessorW const &input, GenericTensorAccessorR const &output) {
  FFK_CHECK(ffk_accumulate(stream, input.ptr, output.ptr, input.shape.num_elements(), ffk_dt(m.data_type)));
}
}  // namespace Reshape

// ------------------------------------------------------------------ concat / split
namespace Concat {
namespace detail {
// blocks = product of dims outside `axis`, block size = product of dims from `axis` inwards
inline void blocks(ArrayShape const &s, ff_dim_t axis, long long &num_blocks, long long &blk) {
  auto d = s.ff_ordered();
  num_blocks = blk = 1;
  for (int i = 0; i < (int)d.size(); ++i) (i < axis.value ? num_blocks : blk) *= d[i];
}
}  // namespace detail
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorW const &output,
                           std::vector<GenericTensorAccessorR> const &inputs, ff_dim_t axis) {
  long long nb, oblk, off = 0;
  detail::blocks(output.shape, axis, nb, oblk);
  size_t es = size_of_datatype(output.data_type);
  for (auto const &in : inputs) {
    long long inb, iblk;
    detail::blocks(in.shape, axis, inb, iblk);
    if (inb != nb) FatalError("Concat: inputs disagree outside the concatenation axis");
    FFK_CHECK(ffk_copy_with_stride(stream, static_cast<char *>(output.ptr) + off * es, in.ptr, nb,
                                   oblk, iblk, ffk_dt(output.data_type)));
    off += iblk;
  }
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorR const &output_grad,
                            std::vector<GenericTensorAccessorW> const &input_grads, ff_dim_t axis) {
  long long nb, oblk, off = 0;
  detail::blocks(output_grad.shape, axis, nb, oblk);
  size_t es = size_of_datatype(output_grad.data_type);
  for (auto const &g : input_grads) {
    long long inb, iblk;
    detail::blocks(g.shape, axis, inb, iblk);
    FFK_CHECK(ffk_add_with_stride(stream, g.ptr, static_cast<char const *>(output_grad.ptr) + off * es,
                                  nb, iblk, oblk, ffk_dt(output_grad.data_type)));
    off += iblk;
  }
}
}  // namespace Concat
namespace Split {
inline void forward_kernel(ffStream_t stream, float **out_ptrs, float const *in_ptr,
                           coord_t const *out_blk_sizes, coord_t in_blk_size, coord_t num_blks,
                           int numOutputs) {
  for (int i = 0; i < numOutputs; i++) {
    FFK_CHECK(ffk_copy_with_stride(stream, out_ptrs[i], in_ptr, num_blks, out_blk_sizes[i], in_blk_size, FFK_FLOAT));
    in_ptr += out_blk_sizes[i];
  }
}
inline void backward_kernel(ffStream_t stream, float *in_grad_ptr, float const **out_grad_ptr,
                            coord_t const *out_blk_sizes, coord_t in_blk_size, coord_t num_blks,
                            int numOutputs) {
  for (int i = 0; i < numOutputs; i++) {
    FFK_CHECK(ffk_add_with_stride(stream, in_grad_ptr, out_grad_ptr[i], num_blks, in_blk_size, out_blk_sizes[i], FFK_FLOAT));
    in_grad_ptr += out_blk_sizes[i];
  }
}
}  // namespace Split

// ------------------------------------------------------------------ transpose / reverse
namespace Transpose {
inline TransposePerDeviceState init_kernel(int num_dim, std::vector<ff_dim_t> const &perm) {
  if ((int)perm.size() != num_dim) FatalError("Transpose: perm must list every dimension");
  return TransposePerDeviceState{num_dim, perm};
}
inline void forward_kernel(cudaStream_t stream, TransposePerDeviceState const &m,
                           GenericTensorAccessorR const &input, GenericTensorAccessorW const &output) {
  auto d = input.shape.ff_ordered();
  std::vector<int> p;
  for (auto x : m.perm) p.push_back(x.value);
  FFK_CHECK(ffk_transpose(stream, input.ptr, output.ptr, m.num_dim, d.data(), p.data(), 0, ffk_dt(input.data_type)));
}
inline void backward_kernel(cudaStream_t stream, TransposePerDeviceState const &m,
                            GenericTensorAccessorW const &in_grad, GenericTensorAccessorR const &out_grad) {
  auto d = out_grad.shape.ff_ordered();
  std::vector<int> inv(m.num_dim);
  for (int i = 0; i < m.num_dim; ++i) inv[m.perm[i].value] = i;
  FFK_CHECK(ffk_transpose(stream, out_grad.ptr, in_grad.ptr, m.num_dim, d.data(), inv.data(), 1, ffk_dt(out_grad.data_type)));
}
}  // namespace Transpose
namespace Reverse {
inline void forward_kernel(ffStream_t stream, float const *in_ptr, float *out_ptr, coord_t num_out_blks,
                           coord_t reverse_dim_size, coord_t in_blk_size, coord_t /*output_size*/) {
  FFK_CHECK(ffk_reverse(stream, in_ptr, out_ptr, num_out_blks, reverse_dim_size, in_blk_size, FFK_FLOAT));
}
inline void backward_kernel(ffStream_t stream, float const *out_grad_ptr, float *in_grad_ptr,
                            coord_t num_out_blks, coord_t reverse_dim_size, coord_t in_blk_size,
                            coord_t /*input_size*/) {
  FFK_CHECK(ffk_reverse(stream, out_grad_ptr, in_grad_ptr, num_out_blks, reverse_dim_size, in_blk_size, FFK_FLOAT));
}
}  // namespace Reverse

// ------------------------------------------------------------------ gather / topk
namespace Gather {
namespace detail {
inline void geom(GenericTensorAccessorR const &big, ArrayShape const &out, legion_dim_t d, long long &n,
                 long long &stride, long long &in_dim, long long &out_dim) {
  stride = 1;
  for (int i = 0; i < d.value; ++i) stride *= (long long)out.dims[i];  // legion dims below `d`
  n = (long long)out.num_elements();
  in_dim = (long long)big.shape[d];
  out_dim = (long long)out[d];
}
}  // namespace detail
inline void forward_kernel(ffStream_t stream, GatherPerDeviceState const &m,
                           GenericTensorAccessorR const &input, GenericTensorAccessorR const &index,
                           GenericTensorAccessorW const &output) {
  long long n, stride, idim, odim;
  detail::geom(input, output.shape, m.legion_dim, n, stride, idim, odim);
  FFK_CHECK(ffk_gather_forward(stream, input.ptr, index.ptr, output.ptr, n, stride, idim, odim,
                               ffk_dt(index.data_type), ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, GatherPerDeviceState const &m,
                            GenericTensorAccessorR const &output_grad, GenericTensorAccessorR const &index,
                            GenericTensorAccessorW const &input_grad) {
  long long n, stride, idim, odim;
  detail::geom(read_only_accessor_from_write_accessor(input_grad), output_grad.shape, m.legion_dim, n,
               stride, idim, odim);
  FFK_CHECK(ffk_gather_backward(stream, output_grad.ptr, index.ptr, input_grad.ptr, n, stride, idim,
                                odim, ffk_dt(index.data_type), ffk_dt(output_grad.data_type)));
}
}  // namespace Gather
namespace TopK {
inline TopKPerDeviceState init_kernel(bool sorted) { return TopKPerDeviceState{sorted}; }
inline void forward_kernel(ffStream_t stream, TopKPerDeviceState const &, float const *input_ptr,
                           float *output_ptr, int *indices_ptr, size_t batch_size, int length, int k,
                           bool sorted) {
  FFK_CHECK(ffk_topk_forward(stream, input_ptr, output_ptr, indices_ptr, batch_size, length, k, sorted));
}
inline void backward_kernel(ffStream_t stream, TopKPerDeviceState const &, float const *out_grad_ptr,
                            int const *indices_ptr, float *in_grad_ptr, size_t batch_size, int length, int k) {
  FFK_CHECK(ffk_topk_backward(stream, out_grad_ptr, indices_ptr, in_grad_ptr, batch_size, length, k));
}
}  // namespace TopK

// ------------------------------------------------------------------ parallel ops (single-device forms)
namespace Replicate {
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                           GenericTensorAccessorW const &output) {
  FFK_CHECK(ffk_copy(stream, output.ptr, input.ptr, input.shape.num_elements(), ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorW const &input,
                            GenericTensorAccessorR const &output, size_t num_replicas) {
  FFK_CHECK(ffk_sum_replicas(stream, input.ptr, output.ptr, input.shape.num_elements(),
                             (int)num_replicas, 1, ffk_dt(input.data_type)));
}
}  // namespace Replicate
namespace Reduction {
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                           GenericTensorAccessorW const &output, size_t num_replicas) {
  FFK_CHECK(ffk_sum_replicas(stream, output.ptr, input.ptr, output.shape.num_elements(),
                             (int)num_replicas, 0, ffk_dt(output.data_type)));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorW const &input,
                            GenericTensorAccessorR const &output) {
  FFK_CHECK(ffk_copy(stream, input.ptr, output.ptr, output.shape.num_elements(), ffk_dt(output.data_type)));
}
}  // namespace Reduction
namespace Combine {
inline void forward_kernel(ffStream_t stream, GenericTensorAccessorR const &input,
                           GenericTensorAccessorW const &output) {
  FFK_CHECK(ffk_copy(stream, output.ptr, input.ptr, input.shape.num_elements(), ffk_dt(input.data_type)));
}
inline void backward_kernel(ffStream_t stream, GenericTensorAccessorR const &output_grad,
                            GenericTensorAccessorW const &input_grad) {
  FFK_CHECK(ffk_accumulate(stream, input_grad.ptr, output_grad.ptr, output_grad.shape.num_elements(),
                           ffk_dt(output_grad.data_type)));
}
}  // namespace Combine
namespace Repartition {
inline RepartitionPerDeviceState init_kernel(PerDeviceFFHandle const &handle, DataType data_type) {
  return RepartitionPerDeviceState{handle, data_type};
}
inline void forward_kernel(ffStream_t stream, RepartitionPerDeviceState const &m,
                           GenericTensorAccessorR const &input, GenericTensorAccessorW const &output) {
  FFK_CHECK(ffk_copy(stream, output.ptr, input.ptr, input.shape.num_elements(), ffk_dt(m.data_type)));
}
// kernel contract pinned by the reference's test (test_partition_kernel.cc:41-58): the first
// (writable) accessor accumulates the second
inline void backward_kernel(ffStream_t stream, RepartitionPerDeviceState const &m,
                            GenericTensorAccessorW const &output_grad, GenericTensorAccessorR const &input_grad) {
  FFK_CHECK(ffk_accumulate(stream, output_grad.ptr, input_grad.ptr, input_grad.shape.num_elements(),
                           ffk_dt(m.data_type)));
}
}  // namespace Repartition

}  // namespace Kernels

// ------------------------------------------------------------------ loss_function_kernels.h
inline void sparse_categorical_crossentropy_loss_backward_kernel(
    ffStream_t stream, float *logit_grad_ptr, float const *logit_ptr, int const *label_ptr,
    size_t /*logit_volume*/, size_t /*logit_grad_volume*/, int num_samples, int num_classes, int k,
    float scale_factor) {
  FFK_CHECK(ffk_sparse_categorical_crossentropy_loss_backward(stream, logit_grad_ptr, logit_ptr, label_ptr,
                                                              num_samples, num_classes, k, scale_factor, FFK_FLOAT));
}
inline void categorical_crossentropy_loss_backward_kernel(ffStream_t stream, float *logit_grad_ptr,
                                                          float const *logit_ptr, float const *label_ptr,
                                                          size_t logit_volume, size_t, float scale_factor) {
  FFK_CHECK(ffk_categorical_crossentropy_loss_backward(stream, logit_grad_ptr, logit_ptr, label_ptr,
                                                       logit_volume, scale_factor, FFK_FLOAT));
}
inline void mean_squared_error_avg_loss_backward_kernel(ffStream_t stream, float *logit_grad_ptr,
                                                        float const *logit_ptr, float const *label_ptr,
                                                        size_t logit_volume, size_t, float scale_factor) {
  FFK_CHECK(ffk_mean_squared_error_avg_loss_backward(stream, logit_grad_ptr, logit_ptr, label_ptr,
                                                     logit_volume, scale_factor, FFK_FLOAT));
}
inline void identity_loss_backward_kernel(ffStream_t stream, float *loss_grad_ptr, float const *loss_ptr,
                                          size_t loss_volume, size_t, float scale_factor) {
  FFK_CHECK(ffk_identity_loss_backward(stream, loss_grad_ptr, loss_ptr, loss_volume, scale_factor, FFK_FLOAT));
}

// ------------------------------------------------------------------ metrics_kernels.h / perf_metrics.h
// LossFunction (op-attrs/ops/loss_functions) and the metric selection the reference keeps beside it
// (lib/runtime/src/metrics_functions.h:28-52).
enum class LossFunction {
  CATEGORICAL_CROSSENTROPY,
  SPARSE_CATEGORICAL_CROSSENTROPY,
  MEAN_SQUARED_ERROR_AVG_REDUCE,
  MEAN_SQUARED_ERROR_SUM_REDUCE,
  IDENTITY
};
enum class Metric {
  ACCURACY,
  CATEGORICAL_CROSSENTROPY,
  SPARSE_CATEGORICAL_CROSSENTROPY,
  MEAN_SQUARED_ERROR,
  ROOT_MEAN_SQUARED_ERROR,
  MEAN_ABSOLUTE_ERROR,
};
class MetricsAttrs {
public:
  MetricsAttrs() = delete;
  MetricsAttrs(LossFunction loss, std::vector<Metric> const &metrics)
      : loss_type(loss), measure_accuracy(false), measure_categorical_crossentropy(false),
        measure_sparse_categorical_crossentropy(false), measure_mean_squared_error(false),
        measure_root_mean_squared_error(false), measure_mean_absolute_error(false) {
    for (Metric m : metrics) {      // lib/runtime/src/metrics_functions.cc:33-58
      switch (m) {
        case Metric::ACCURACY: measure_accuracy = true; break;
        case Metric::CATEGORICAL_CROSSENTROPY: measure_categorical_crossentropy = true; break;
        case Metric::SPARSE_CATEGORICAL_CROSSENTROPY: measure_sparse_categorical_crossentropy = true; break;
        case Metric::MEAN_SQUARED_ERROR: measure_mean_squared_error = true; break;
        case Metric::ROOT_MEAN_SQUARED_ERROR: measure_root_mean_squared_error = true; break;
        case Metric::MEAN_ABSOLUTE_ERROR: measure_mean_absolute_error = true; break;
      }
    }
  }
  LossFunction loss_type;
  bool measure_accuracy;
  bool measure_categorical_crossentropy;
  bool measure_sparse_categorical_crossentropy;
  bool measure_mean_squared_error;
  bool measure_root_mean_squared_error;
  bool measure_mean_absolute_error;
};
inline int ffk_metric_flags(MetricsAttrs const &m) {
  return (m.measure_accuracy ? FFK_METRIC_ACCURACY : 0) | (m.measure_categorical_crossentropy ? FFK_METRIC_CCE : 0) |
         (m.measure_sparse_categorical_crossentropy ? FFK_METRIC_SPARSE_CCE : 0) |
         (m.measure_mean_squared_error ? FFK_METRIC_MSE : 0) |
         (m.measure_root_mean_squared_error ? FFK_METRIC_RMSE : 0) |
         (m.measure_mean_absolute_error ? FFK_METRIC_MAE : 0);
}
// src/perf_metrics.cc:29-68
inline PerfMetrics update(PerfMetrics const &lhs, PerfMetrics const &rhs) {
  PerfMetrics out(lhs);
  auto add = [](std::optional<float> &l, std::optional<float> const &r) {
    if (l.has_value()) l.value() += r.value_or(0.f);
  };
  out.train_all += rhs.train_all;
  if (out.train_correct.has_value()) out.train_correct.value() += rhs.train_correct.value_or(0);
  add(out.cce_loss, rhs.cce_loss);
  add(out.sparse_cce_loss, rhs.sparse_cce_loss);
  add(out.mse_loss, rhs.mse_loss);
  add(out.rmse_loss, rhs.rmse_loss);
  add(out.mae_loss, rhs.mae_loss);
  out.current_time = rhs.current_time;
  return out;
}
inline PerfMetrics apply_scale(PerfMetrics const &pm, float scale) {
  PerfMetrics out(pm);
  for (std::optional<float> *l : {&out.cce_loss, &out.sparse_cce_loss, &out.mse_loss, &out.rmse_loss, &out.mae_loss})
    if (l->has_value()) l->value() *= scale;
  return out;
}
// The reference's kernels add into a zero-copy (host-mapped) PerfMetrics (metrics_functions.cu:23-137).
// Here the batch is accumulated in a small device record (one per host thread and device, kept), then
// read back and added to `perf_zc` — the call returns with `perf_zc` up to date, as a caller that
// reads its zero-copy record after synchronising would see it.
namespace detail {
inline ffk_perf_metrics_t *metrics_scratch() {
  thread_local ffk_perf_metrics_t *buf[64] = {};
  int dev = 0;
  checkCUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) dev = 0;
  if (!buf[dev]) checkCUDA(cudaMalloc(reinterpret_cast<void **>(&buf[dev]), sizeof(ffk_perf_metrics_t)));
  return buf[dev];
}
inline void metrics_fold(ffStream_t stream, ffk_perf_metrics_t const *dev_rec, PerfMetrics &perf_zc) {
  ffk_perf_metrics_t h;
  checkCUDA(cudaMemcpyAsync(&h, dev_rec, sizeof(h), cudaMemcpyDeviceToHost, stream));
  checkCUDA(cudaStreamSynchronize(stream));
  perf_zc.train_all += h.train_all;
  if (perf_zc.train_correct.has_value()) perf_zc.train_correct.value() += h.train_correct;
  if (perf_zc.cce_loss.has_value()) perf_zc.cce_loss.value() += h.cce_loss;
  if (perf_zc.sparse_cce_loss.has_value()) perf_zc.sparse_cce_loss.value() += h.sparse_cce_loss;
  if (perf_zc.mse_loss.has_value()) perf_zc.mse_loss.value() += h.mse_loss;
  if (perf_zc.rmse_loss.has_value()) perf_zc.rmse_loss.value() += h.rmse_loss;
  if (perf_zc.mae_loss.has_value()) perf_zc.mae_loss.value() += h.mae_loss;
}
}  // namespace detail
inline void update_metrics_sparse_label_kernel(ffStream_t stream, MetricsAttrs const &metrics, float const *logit_ptr,
                                               int const *label_ptr, int num_samples, int num_classes,
                                               PerfMetrics &perf_zc) {
  ffk_perf_metrics_t *rec = detail::metrics_scratch();
  checkCUDA(cudaMemsetAsync(rec, 0, sizeof(*rec), stream));
  FFK_CHECK(ffk_update_metrics_sparse_label(stream, ffk_metric_flags(metrics), logit_ptr, label_ptr, num_samples,
                                            num_classes, rec));
  detail::metrics_fold(stream, rec, perf_zc);
}
inline void update_metrics_label_kernel(ffStream_t stream, MetricsAttrs const &metrics, float const *logit_ptr,
                                        float const *label_ptr, int num_samples, int num_classes,
                                        PerfMetrics &perf_zc) {
  ffk_perf_metrics_t *rec = detail::metrics_scratch();
  checkCUDA(cudaMemsetAsync(rec, 0, sizeof(*rec), stream));
  FFK_CHECK(ffk_update_metrics_label(stream, ffk_metric_flags(metrics), logit_ptr, label_ptr, num_samples,
                                     num_classes, rec));
  detail::metrics_fold(stream, rec, perf_zc);
}

// ------------------------------------------------------------------ initializer_kernels.h, cpu.h
// zero / constant initialisation (initializer_kernels.h:21-28). The reference has the CPU bodies only
// (src/cpu/initializer_kernels.cc:1-52; its zero_init_kernel_gpu is NOT_IMPLEMENTED, constant_init_kernel
// and the three random initialisers are declared and defined nowhere in lib/kernels). The GPU forms here
// are ffk_fill on the calling thread's stream; the random initialisers follow below.
enum class TaskLocation { CPU, GPU };
struct DataTypeValue {   // op-attrs/datatype_value.dtg.h: one value of any tensor element type
  std::variant<float, double, int32_t, int64_t, bool> raw;
  template <typename T> DataTypeValue(T v) : raw(v) {}
  template <typename T> bool has() const { return std::holds_alternative<T>(raw); }
  template <typename T> T get() const { return std::get<T>(raw); }
  double as_double() const {
    return std::visit([](auto v) { return static_cast<double>(v); }, raw);
  }
};
inline void constant_init_kernel_gpu(GenericTensorAccessorW const &tensor, DataTypeValue value) {
  ffStream_t stream;
  checkCUDA(get_legion_stream(&stream));
  FFK_CHECK(ffk_fill(stream, tensor.ptr, tensor.shape.get_volume(), value.as_double(), ffk_dt(tensor.data_type)));
  checkCUDA(cudaStreamSynchronize(stream));
}
inline void constant_init_kernel_cpu(GenericTensorAccessorW const &tensor, DataTypeValue value) {
  size_t const n = tensor.shape.get_volume();
  double const v = value.as_double();
  switch (tensor.data_type) {
    case DataType::FLOAT: for (size_t i = 0; i < n; ++i) static_cast<float *>(tensor.ptr)[i] = float(v); break;
    case DataType::DOUBLE: for (size_t i = 0; i < n; ++i) static_cast<double *>(tensor.ptr)[i] = v; break;
    case DataType::INT32: for (size_t i = 0; i < n; ++i) static_cast<int32_t *>(tensor.ptr)[i] = int32_t(v); break;
    case DataType::INT64: for (size_t i = 0; i < n; ++i) static_cast<int64_t *>(tensor.ptr)[i] = int64_t(v); break;
    case DataType::BOOL: for (size_t i = 0; i < n; ++i) static_cast<bool *>(tensor.ptr)[i] = v != 0.0; break;
    default: throw std::runtime_error("constant_init_kernel_cpu: 16-bit tensors have no host element type here");
  }
}
inline void constant_init_kernel(TaskLocation const &loc, GenericTensorAccessorW const &tensor, DataTypeValue value) {
  if (loc == TaskLocation::CPU) constant_init_kernel_cpu(tensor, value);
  else constant_init_kernel_gpu(tensor, value);
}
inline void zero_init_kernel_gpu(GenericTensorAccessorW const &tensor) { constant_init_kernel_gpu(tensor, DataTypeValue(0.f)); }
inline void zero_init_kernel_cpu(GenericTensorAccessorW const &tensor) {
  std::memset(tensor.ptr, 0, tensor.shape.get_volume() * size_of_datatype(tensor.data_type));
}
inline void zero_init_kernel(TaskLocation const &loc, GenericTensorAccessorW const &tensor) {
  if (loc == TaskLocation::CPU) zero_init_kernel_cpu(tensor);
  else zero_init_kernel_gpu(tensor);
}

// The three random initialisers (initializer_kernels.h:10-20). lib/kernels declares them and defines them
// nowhere; the callers (lib/runtime/src/initializer.cc:88-108, 203-229) fix the meaning: uniform in
// [min_val, max_val]; Glorot = uniform in [-scale, scale] with the caller's scale = sqrt(6 / (fan_in +
// fan_out)); normal(mean, stddev); float tensors only (the legacy cuRAND bodies assert DT_FLOAT,
// lib/runtime/src/initializer_kernels.cu:35). Initialisation runs once, off the hot path: the values are a
// function of the seed alone (SplitMix64 counter stream, 24-bit mantissas, Box-Muller), generated on the
// host and copied to wherever the tensor lives.
namespace detail {
inline uint64_t splitmix64(uint64_t &state) {
  uint64_t z = (state += 0x9E3779B97F4A7C15ULL);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
inline float unit_float(uint64_t &state) { return float(splitmix64(state) >> 40) * (1.0f / 16777216.0f); }  // [0, 1)
inline void require_float(GenericTensorAccessorW const &t, char const *who) {
  if (t.data_type != DataType::FLOAT) throw mk_runtime_error(std::string(who) + ": float tensors only");
}
inline void store_floats(GenericTensorAccessorW const &t, std::vector<float> const &host) {
  cudaPointerAttributes attr{};
  bool const on_device = cudaPointerGetAttributes(&attr, t.ptr) == cudaSuccess &&
                         (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
  if (!on_device) {
    cudaGetLastError();  // a plain host pointer (or no driver at all) is not an error here
    std::memcpy(t.ptr, host.data(), host.size() * sizeof(float));
    return;
  }
  checkCUDA(cudaMemcpy(t.ptr, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice));
}
}  // namespace detail
inline void uniform_init_kernel(GenericTensorAccessorW const &tensor, size_t seed, float min_val, float max_val) {
  detail::require_float(tensor, "uniform_init_kernel");
  std::vector<float> host(tensor.shape.get_volume());
  uint64_t state = seed;
  for (float &v : host) v = min_val + (max_val - min_val) * detail::unit_float(state);
  detail::store_floats(tensor, host);
}
inline void glorot_uniform_init_kernel(GenericTensorAccessorW const &tensor, size_t seed, float scale_factor) {
  uniform_init_kernel(tensor, seed, -scale_factor, scale_factor);
}
inline void norm_init_kernel(GenericTensorAccessorW const &tensor, size_t seed, float mean, float stddev) {
  detail::require_float(tensor, "norm_init_kernel");
  std::vector<float> host(tensor.shape.get_volume());
  uint64_t state = seed;
  for (size_t i = 0; i < host.size(); i += 2) {
    float const u1 = 1.0f - detail::unit_float(state), u2 = detail::unit_float(state);   // u1 in (0, 1]
    float const r = std::sqrt(-2.0f * std::log(u1)), a = 6.28318530717958647692f * u2;
    host[i] = mean + stddev * r * std::cos(a);
    if (i + 1 < host.size()) host[i + 1] = mean + stddev * r * std::sin(a);
  }
  detail::store_floats(tensor, host);
}

// ------------------------------------------------------------------ optimizer_kernels.h
// (the reference header has a syntax error at :23 — a missing comma — fixed here)
inline void sgd_ps_update_task_gpu(ffStream_t stream, float lr, float momentum, bool nesterov,
                                   float weight_decay, float const *weight_grad_ptr, size_t size,
                                   int num_replicas, float *weight_ptr, float *sgd_v_ptr) {
  FFK_CHECK(ffk_sgd_update(stream, lr, momentum, nesterov, weight_decay, weight_grad_ptr, size,
                           num_replicas, weight_ptr, sgd_v_ptr, FFK_FLOAT, nullptr));
}
inline void adam_ps_update_task_gpu(ffStream_t stream, float alpha_t, float beta1, float beta2,
                                    float weight_decay, float epsilon, float const *weight_grad_ptr,
                                    size_t size, int num_replicas, float *adam_m_ptr, float *adam_v_ptr,
                                    float *weight_ptr) {
  FFK_CHECK(ffk_adam_update(stream, alpha_t, beta1, beta2, weight_decay, epsilon, weight_grad_ptr, size,
                            num_replicas, adam_m_ptr, adam_v_ptr, weight_ptr, FFK_FLOAT, nullptr));
}

// the *_nccl variants: all-reduce (SUM, no division) on handle.ncclComm, then the same update
inline void sgd_nccl_update_task_gpu(ffStream_t stream, float lr, float momentum, bool nesterov,
                                     float weight_decay, PerDeviceFFHandle const &handle,
                                     float const *weight_grad_ptr, size_t size, float *weight_ptr,
                                     float *sgd_v_ptr) {
  FFK_CHECK(ffk_sgd_nccl_update(stream, handle.ncclComm, lr, momentum, nesterov, weight_decay,
                                const_cast<float *>(weight_grad_ptr), size, weight_ptr, sgd_v_ptr,
                                FFK_FLOAT, nullptr));
}
inline void adam_nccl_update_task_gpu(ffStream_t stream, float alpha_t, float beta1, float beta2,
                                      float weight_decay, float epsilon, PerDeviceFFHandle const &handle,
                                      float const *weight_grad_ptr, size_t size, float *adam_m_ptr,
                                      float *adam_v_ptr, float *weight_ptr) {
  FFK_CHECK(ffk_adam_nccl_update(stream, handle.ncclComm, alpha_t, beta1, beta2, weight_decay, epsilon,
                                 const_cast<float *>(weight_grad_ptr), size, adam_m_ptr, adam_v_ptr,
                                 weight_ptr, FFK_FLOAT, nullptr));
}

// ------------------------------------------------------------------ nccl.h
struct ncclUniqueIdBytes {
  char internal[FFK_NCCL_UNIQUE_ID_BYTES];
};
namespace Kernels {
namespace NCCL {
inline ncclUniqueIdBytes generate_unique_id() {
  ncclUniqueIdBytes id;
  FFK_CHECK(ffk_nccl_get_unique_id(id.internal));
  return id;
}
inline void *create_comm(ncclUniqueIdBytes const &id, int num_ranks, int my_rank) {
  void *comm = nullptr;
  FFK_CHECK(ffk_nccl_comm_create(&comm, id.internal, num_ranks, my_rank));
  return comm;
}
}  // namespace NCCL
}  // namespace Kernels

}  // namespace FlexFlow
