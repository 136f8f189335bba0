// C++ veneer, part 1: the value types the reference's kernel API is written in
// (SURVEY App. B), hand-written so that callers compile without the reference's generated
// *.dtg.h headers. Names, enumerator order and member names follow the reference; the
// implementations are this repository's.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <unordered_set>
#include <utility>
#include <variant>
#include <vector>

#include "ffk.h"

#ifndef MAX_TENSOR_DIM
#define MAX_TENSOR_DIM 5
#endif

typedef cudaStream_t ffStream_t;
typedef cudaEvent_t ffEvent_t;
typedef cudaError_t ffError_t;
// The reference types these as cuDNN / cuBLAS handles (include/kernels/device.h:33-55).
// Nothing here calls those libraries; the slots survive as opaque pointers.
typedef void *ffHandle_t;
typedef void *ffblasHandle_t;

namespace FlexFlow {

using coord_t = long long;

// ---- error conventions (include/kernels/device.h:88-108, utils/exception.h)
[[noreturn]] inline void fatal_error(std::string const &msg, char const *file, int line) {
  std::fprintf(stderr, "%s\n%s:%d\nAborting...\n", msg.c_str(), file, line);
  std::exit(1);
}
#define FatalError(s) ::FlexFlow::fatal_error((s), __FILE__, __LINE__)
#define checkCUDA(status)                                                                        \
  do {                                                                                           \
    cudaError_t _st = (status);                                                                  \
    if (_st != cudaSuccess) {                                                                    \
      FatalError(std::string("CUDA failure: ") + cudaGetErrorString(_st) + " (" +               \
                 std::to_string((int)_st) + ")");                                                \
    }                                                                                            \
  } while (0)
// C-ABI status -> the reference's behaviour on a failed kernel call: print and abort.
#define FFK_CHECK(call)                                                                          \
  do {                                                                                           \
    if ((call) != FFK_OK) { FatalError(std::string("ffk failure: ") + ffk_last_error()); }       \
  } while (0)

// device.h:86,110-114 — the event helpers of profiling_wrapper and the error string
inline char const *getCudaErrorString(cudaError_t status) { return cudaGetErrorString(status); }
inline ffError_t ffEventCreate(ffEvent_t *e) { return cudaEventCreate(e); }
inline ffError_t ffEventDestroy(ffEvent_t &e) { return cudaEventDestroy(e); }
inline ffError_t ffEventRecord(ffEvent_t &e, ffStream_t stream) { return cudaEventRecord(e, stream); }
inline ffError_t ffEventSynchronize(ffEvent_t &e) { return cudaEventSynchronize(e); }
inline ffError_t ffEventElapsedTime(float *elapsed, ffEvent_t &start, ffEvent_t &stop) {
  return cudaEventElapsedTime(elapsed, start, stop);
}

inline std::runtime_error mk_runtime_error(std::string const &s) { return std::runtime_error(s); }
#define NOT_IMPLEMENTED() throw ::FlexFlow::mk_runtime_error("not implemented")

// ---- enums (order = the reference's .enum.toml files)
enum class DataType { BOOL, INT32, INT64, HALF, FLOAT, DOUBLE, BFLOAT16 /* extension */ };

inline size_t size_of_datatype(DataType dt) {
  switch (dt) {
    case DataType::BOOL: return sizeof(bool);
    case DataType::INT32: return 4;
    case DataType::INT64: return 8;
    case DataType::HALF: return 2;
    case DataType::FLOAT: return 4;
    case DataType::DOUBLE: return 8;
    case DataType::BFLOAT16: return 2;
  }
  throw mk_runtime_error("unknown DataType");
}
inline char const *format_as(DataType dt) {
  static char const *names[] = {"BOOL", "INT32", "INT64", "HALF", "FLOAT", "DOUBLE", "BFLOAT16"};
  return names[static_cast<int>(dt)];
}

struct bfloat16_tag { uint16_t bits; };  // host-side placeholder for 2-byte bf16 storage

template <DataType> struct data_type_enum_to_class;
template <> struct data_type_enum_to_class<DataType::BOOL> { using type = bool; };
template <> struct data_type_enum_to_class<DataType::INT32> { using type = int32_t; };
template <> struct data_type_enum_to_class<DataType::INT64> { using type = int64_t; };
template <> struct data_type_enum_to_class<DataType::HALF> { using type = half; };
template <> struct data_type_enum_to_class<DataType::FLOAT> { using type = float; };
template <> struct data_type_enum_to_class<DataType::DOUBLE> { using type = double; };
template <> struct data_type_enum_to_class<DataType::BFLOAT16> { using type = bfloat16_tag; };
template <DataType DT> using real_type_t = typename data_type_enum_to_class<DT>::type;

enum class OperatorType {
  NOOP, INPUT, WEIGHT, CONV2D, DROPOUT, LINEAR, BATCHMATMUL, POOL2D, SCALAR_MULTIPLY, SCALAR_ADD,
  SCALAR_FLOOR_DIV, SCALAR_TRUE_DIV, SCALAR_SUB, RELU, IDENTITY, SIGMOID, TANH, ELU, FLAT, SOFTMAX,
  BATCHNORM, CONCAT, SPLIT, EMBEDDING, CACHE, RESHAPE, REVERSE, TRANSPOSE, EW_ADD, EW_MUL, MATMUL,
  MUL, ENLARGE, SQUEEZE, UNSQUEEZE, EW_SUB, EW_DIV, EW_EQUAL, EW_GREATER, EW_LESS, EW_MAX, EW_MIN,
  REDUCE_ARGMAX, REDUCE_ARGMIN, REDUCE_MAX, REDUCE_MEAN, REDUCE_MIN, REDUCE_PROD, REDUCE_SUM, PAD,
  SHAPE, SIZE, TOPK, WHERE, CEIL, CAST, EXP, ROUND, LOG, LOGICAL_NOT, SQRT, SIN, COS, LEAKYRELU,
  SLICE, RESIZE, PRELU, GELU, MULTIHEAD_ATTENTION, FUSED, RSQRT, POW, MEAN, LAYERNORM, GATHER,
  BROADCAST, REPARTITION, COMBINE, REPLICATE, REDUCTION, BATCH, PIPELINE, FUSED_PARALLEL
};
static_assert(static_cast<int>(OperatorType::EW_ADD) == FFK_OP_EW_ADD, "enum order drifted");
static_assert(static_cast<int>(OperatorType::GELU) == FFK_OP_GELU, "enum order drifted");
static_assert(static_cast<int>(OperatorType::POW) == FFK_OP_POW, "enum order drifted");
static_assert(static_cast<int>(OperatorType::REDUCE_SUM) == FFK_OP_REDUCE_SUM, "enum order drifted");
static_assert(static_cast<int>(OperatorType::FUSED_PARALLEL) == 82, "enum order drifted");

enum class Activation { RELU, SIGMOID, TANH, GELU };
enum class AggregateOp { SUM, AVG };
enum class PoolOp { MAX, AVG };

inline int act_code(std::optional<Activation> const &a) {
  return a.has_value() ? static_cast<int>(*a) : FFK_ACT_NONE;
}

// ---- dimension indices
struct ff_dim_t {
  int value;
  constexpr ff_dim_t(int v = 0) : value(v) {}
  bool operator==(ff_dim_t o) const { return value == o.value; }
};
struct legion_dim_t {   // kernels/legion_dim_t.struct.toml: eq, ord, hash
  int value;
  constexpr legion_dim_t(int v = 0) : value(v) {}
  bool operator==(legion_dim_t o) const { return value == o.value; }
  bool operator!=(legion_dim_t o) const { return value != o.value; }
  bool operator<(legion_dim_t o) const { return value < o.value; }
  bool operator>(legion_dim_t o) const { return value > o.value; }
  bool operator<=(legion_dim_t o) const { return value <= o.value; }
  bool operator>=(legion_dim_t o) const { return value >= o.value; }
};
using LegionTensorDims = std::vector<size_t>;  // what ArrayShape::dims holds (array_shape.h:50)
inline legion_dim_t add_to_legion_dim(legion_dim_t d, int v) { return legion_dim_t{d.value + v}; }
inline legion_dim_t legion_dim_from_ff_dim(ff_dim_t d, int num_dims) {
  return legion_dim_t{num_dims - d.value - 1};
}

// ---- tensor shapes
struct TensorDims {
  std::vector<size_t> ff_ordered;  // outermost first
};
struct TensorShape {
  TensorDims dims;
  DataType data_type;
};
inline size_t num_dims(TensorShape const &s) { return s.dims.ff_ordered.size(); }
inline size_t get_num_elements(TensorShape const &s) {
  size_t n = 1;
  for (size_t d : s.dims.ff_ordered) n *= d;
  return n;
}
inline size_t get_size_in_bytes(TensorShape const &s) {
  return get_num_elements(s) * size_of_datatype(s.data_type);
}
// the reference's test helper: dims given innermost-first (legion order)
inline TensorShape make_tensor_shape_from_legion_dims(std::vector<size_t> legion, DataType dt) {
  TensorShape s{{std::vector<size_t>(legion.rbegin(), legion.rend())}, dt};
  return s;
}
#ifndef FFK_NO_TEST_SHAPE_HELPERS  // the reference's kernel tests bring their own (test/src/test_utils.cc)
inline TensorShape make_float_tensor_shape_from_legion_dims(std::vector<size_t> legion) {
  return make_tensor_shape_from_legion_dims(std::move(legion), DataType::FLOAT);
}
#endif

// ArrayShape — include/kernels/array_shape.h:14-52. Dims are held in legion order
// (innermost first); rank-0 shapes have 0 elements, as in the reference.
struct ArrayShape {
  ArrayShape() = delete;
  ArrayShape(size_t const *dims_, size_t n) : dims(dims_, dims_ + n) {}
  explicit ArrayShape(std::vector<size_t> const &legion_ordered) : dims(legion_ordered) {}
  ArrayShape(TensorShape const &s) : dims(s.dims.ff_ordered.rbegin(), s.dims.ff_ordered.rend()) {}

  size_t num_dims() const { return dims.size(); }
  size_t get_dim() const { return num_dims(); }
  size_t num_elements() const {
    if (dims.empty()) return 0;
    size_t n = 1;
    for (size_t d : dims) n *= d;
    return n;
  }
  size_t get_volume() const { return num_elements(); }
  size_t operator[](legion_dim_t i) const { return dims.at(i.value); }
  size_t at(legion_dim_t i) const { return dims.at(i.value); }
  size_t at(ff_dim_t i) const { return dims.at(dims.size() - 1 - i.value); }
  legion_dim_t last_idx() const { return legion_dim_t{(int)dims.size() - 1}; }
  legion_dim_t neg_idx(int i) const { return legion_dim_t{(int)dims.size() + i}; }
  std::optional<size_t> at_maybe(legion_dim_t i) const {
    if (i.value >= 0 && (size_t)i.value < dims.size()) return dims[i.value];
    return std::nullopt;
  }
  std::optional<size_t> at_maybe(ff_dim_t i) const {
    if (i.value >= 0 && (size_t)i.value < dims.size()) return at(i);
    return std::nullopt;
  }
  // Half-open range [start, end) of legion dims (NOT_IMPLEMENTED in the reference,
  // src/array_shape.cc:53-57; Concat and Gather depend on it).
  ArrayShape sub_shape(std::optional<std::variant<ff_dim_t, legion_dim_t>> start,
                       std::optional<std::variant<ff_dim_t, legion_dim_t>> end) const {
    auto to_legion = [&](std::variant<ff_dim_t, legion_dim_t> const &v) -> int {
      if (std::holds_alternative<legion_dim_t>(v)) return std::get<legion_dim_t>(v).value;
      return (int)dims.size() - 1 - std::get<ff_dim_t>(v).value;
    };
    int lo = start ? to_legion(*start) : 0;
    int hi = end ? to_legion(*end) : (int)dims.size();
    if (lo > hi) std::swap(lo, hi);
    lo = std::max(lo, 0);
    hi = std::min(hi, (int)dims.size());
    return ArrayShape(std::vector<size_t>(dims.begin() + lo, dims.begin() + hi));
  }
  std::vector<long long> ff_ordered() const {
    return std::vector<long long>(dims.rbegin(), dims.rend());
  }
  bool operator==(ArrayShape const &o) const { return dims == o.dims; }
  bool operator!=(ArrayShape const &o) const { return dims != o.dims; }

  std::vector<size_t> dims;  // legion order
};
inline size_t get_volume(ArrayShape const &s) { return s.get_volume(); }
inline TensorShape get_tensor_shape(ArrayShape const &s, DataType dt) {
  return TensorShape{{std::vector<size_t>(s.dims.rbegin(), s.dims.rend())}, dt};
}

// ---- accessors — include/kernels/accessor.h:13-66
template <typename P> struct AccessorBase {
  DataType data_type;
  ArrayShape shape;
  P ptr;
  template <DataType DT> auto get() const {
    using T = std::conditional_t<std::is_const_v<std::remove_pointer_t<P>>, real_type_t<DT> const,
                                 real_type_t<DT>>;
    if (data_type != DT)
      throw mk_runtime_error(std::string("Invalid access data type (") + format_as(data_type) +
                             " != " + format_as(DT) + ")");
    return static_cast<T *>(ptr);
  }
  auto get_int32_ptr() const { return get<DataType::INT32>(); }
  auto get_int64_ptr() const { return get<DataType::INT64>(); }
  auto get_float_ptr() const { return get<DataType::FLOAT>(); }
  auto get_double_ptr() const { return get<DataType::DOUBLE>(); }
  auto get_half_ptr() const { return get<DataType::HALF>(); }
};
struct GenericTensorAccessorW : AccessorBase<void *> {};
struct GenericTensorAccessorR : AccessorBase<void const *> {};
inline GenericTensorAccessorR read_only_accessor_from_write_accessor(GenericTensorAccessorW const &w) {
  return GenericTensorAccessorR{{w.data_type, w.shape, w.ptr}};
}
inline bool is_shape_and_dtype_equal(GenericTensorAccessorW const &a, GenericTensorAccessorW const &b) {
  return a.shape == b.shape && a.data_type == b.data_type;
}
template <typename A> bool shape_and_dtype_matches(A const &a, ArrayShape const &s, DataType const &dt) {
  return a.shape == s && a.data_type == dt;
}
inline float *get_float_ptr(GenericTensorAccessorW const &a) { return a.get_float_ptr(); }
inline float const *get_float_ptr(GenericTensorAccessorR const &a) { return a.get_float_ptr(); }
// accessor.h:68-167 — the free-function forms: typed pointer of one accessor (throws on a data type
// mismatch) and of a vector of accessors, per element type
template <DataType DT> real_type_t<DT> *get(GenericTensorAccessorW const &a) { return a.template get<DT>(); }
template <DataType DT> real_type_t<DT> const *get(GenericTensorAccessorR const &a) { return a.template get<DT>(); }
template <DataType DT> std::vector<real_type_t<DT> *> get(std::vector<GenericTensorAccessorW> const &accs) {
  std::vector<real_type_t<DT> *> out;
  for (auto const &a : accs) out.push_back(get<DT>(a));
  return out;
}
template <DataType DT> std::vector<real_type_t<DT> const *> get(std::vector<GenericTensorAccessorR> const &accs) {
  std::vector<real_type_t<DT> const *> out;
  for (auto const &a : accs) out.push_back(get<DT>(a));
  return out;
}
#define FFK_ACCESSOR_PTRS(name, DT)                                                                               \
  inline real_type_t<DT> *get_##name##_ptr(GenericTensorAccessorW const &a) { return get<DT>(a); }                  \
  inline real_type_t<DT> const *get_##name##_ptr(GenericTensorAccessorR const &a) { return get<DT>(a); }            \
  inline std::vector<real_type_t<DT> *> get_##name##_ptrs(std::vector<GenericTensorAccessorW> const &a) {           \
    return get<DT>(a);                                                                                              \
  }                                                                                                                 \
  inline std::vector<real_type_t<DT> const *> get_##name##_ptrs(std::vector<GenericTensorAccessorR> const &a) {     \
    return get<DT>(a);                                                                                              \
  }
FFK_ACCESSOR_PTRS(int32, DataType::INT32)
FFK_ACCESSOR_PTRS(int64, DataType::INT64)
FFK_ACCESSOR_PTRS(double, DataType::DOUBLE)
FFK_ACCESSOR_PTRS(half, DataType::HALF)
#undef FFK_ACCESSOR_PTRS
inline std::vector<float *> get_float_ptrs(std::vector<GenericTensorAccessorW> const &a) { return get<DataType::FLOAT>(a); }
inline std::vector<float const *> get_float_ptrs(std::vector<GenericTensorAccessorR> const &a) {
  return get<DataType::FLOAT>(a);
}
inline std::pair<ArrayShape, DataType> get_shape_and_datatype(GenericTensorAccessorR const &a) {
  return {a.shape, a.data_type};
}
inline std::pair<ArrayShape, DataType> get_shape_and_datatype(GenericTensorAccessorW const &a) {
  return {a.shape, a.data_type};
}

// ---- datatype_dispatch.h:8-88 — run F<DT> (or F<IT, OT>) for data types known at run time
template <template <DataType> class F, typename... Args>
auto dispatch(DataType dt, Args &&...args) -> decltype(F<DataType::FLOAT>{}(std::forward<Args>(args)...)) {
  switch (dt) {
    case DataType::FLOAT: return F<DataType::FLOAT>{}(std::forward<Args>(args)...);
    case DataType::DOUBLE: return F<DataType::DOUBLE>{}(std::forward<Args>(args)...);
    case DataType::INT32: return F<DataType::INT32>{}(std::forward<Args>(args)...);
    case DataType::INT64: return F<DataType::INT64>{}(std::forward<Args>(args)...);
    case DataType::BOOL: return F<DataType::BOOL>{}(std::forward<Args>(args)...);
    default: throw mk_runtime_error(std::string("Unknown datatype ") + format_as(dt));
  }
}
template <template <DataType> class F> struct DataTypeDispatch1 {
  template <typename... Args>
  auto operator()(DataType data_type, Args... args) -> decltype(F<DataType::FLOAT>{}(args...)) {
    return dispatch<F>(data_type, args...);
  }
};
template <template <DataType, DataType> class F> struct DataTypeDispatch2 {
  template <DataType IT> struct InputType {
    template <DataType OT> struct OutputType {
      template <typename... Args> void operator()(Args... args) const { F<IT, OT>{}(args...); }
    };
    template <typename... Args> void operator()(DataType output_type, Args... args) const {
      dispatch<OutputType>(output_type, args...);
    }
  };
  template <typename... Args> void operator()(DataType input_data_type, DataType output_data_type, Args... args) {
    dispatch<InputType>(input_data_type, output_data_type, args...);
  }
};

// ---- attrs consumed by kernels (op-attrs)
struct L1RegularizerAttrs { float lambda; };
struct L2RegularizerAttrs { float lambda; };
struct RegularizerAttrs {
  std::variant<L1RegularizerAttrs, L2RegularizerAttrs> raw;
  RegularizerAttrs(L1RegularizerAttrs a) : raw(a) {}
  RegularizerAttrs(L2RegularizerAttrs a) : raw(a) {}
  template <typename T> bool has() const { return std::holds_alternative<T>(raw); }
  template <typename T> T const &get() const { return std::get<T>(raw); }
};
struct ElementUnaryAttrs {
  OperatorType op_type;
  std::optional<float> scalar;
};
inline OperatorType get_op_type(ElementUnaryAttrs const &a) { return a.op_type; }
struct ProfilingSettings {   // kernels/profiling_settings.struct.toml: eq, ord, hash
  int warmup_iters;
  int measure_iters;
  bool operator==(ProfilingSettings const &o) const {
    return warmup_iters == o.warmup_iters && measure_iters == o.measure_iters;
  }
  bool operator!=(ProfilingSettings const &o) const { return !(*this == o); }
  bool operator<(ProfilingSettings const &o) const {
    return warmup_iters != o.warmup_iters ? warmup_iters < o.warmup_iters : measure_iters < o.measure_iters;
  }
};

// ---- PerDeviceFFHandle — include/kernels/ff_handle.h:13-24
struct PerDeviceFFHandle {
  ffHandle_t dnn;
  ffblasHandle_t blas;
  void *workSpace;
  size_t workSpaceSize;
  bool allowTensorOpMathConversion;
  void *ncclComm;
};

}  // namespace FlexFlow

namespace std {
template <> struct hash<::FlexFlow::legion_dim_t> {
  size_t operator()(::FlexFlow::legion_dim_t const &d) const { return std::hash<int>{}(d.value); }
};
template <> struct hash<::FlexFlow::ProfilingSettings> {
  size_t operator()(::FlexFlow::ProfilingSettings const &p) const {
    return std::hash<int>{}(p.warmup_iters) * 1000003u ^ std::hash<int>{}(p.measure_iters);
  }
};
}  // namespace std
