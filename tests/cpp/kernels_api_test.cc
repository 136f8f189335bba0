// The reference's kernel tests (lib/kernels/test/src/test_*.cc), re-expressed against this
// repository's veneer of the same API: same shapes, same fills, same assertions ("output
// contains a non-zero" where the reference asserts only that; exact vectors where it pins
// them), plus the arena / profiling plumbing. Built by __graft_entry__.build(), run on the GPU
// by tests/test_cpp_api_gpu.py. No test framework: a CHECK macro and a case counter.
#include <cmath>
#include <cstdio>
#include <functional>
#include <random>
#include <vector>

#include "kernels/accessor.h"
#include "kernels/linear_kernels.h"

using namespace FlexFlow;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                                      \
  do {                                                                                   \
    ++g_checks;                                                                          \
    if (!(cond)) {                                                                       \
      ++g_fail;                                                                          \
      std::printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                \
    }                                                                                    \
  } while (0)

// ---- helpers mirroring lib/kernels/test/src/test_utils.{h,cc}
static std::mt19937 rng(1234);  // seeded (the reference's random_device fill is not reproducible)
template <typename T> std::vector<T> to_host(void const *dev, size_t n) {
  std::vector<T> h(n);
  checkCUDA(cudaMemcpy(h.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost));
  return h;
}
static GenericTensorAccessorW filled(TensorShape const &s, Allocator &a, std::function<float(size_t)> f) {
  GenericTensorAccessorW acc = a.allocate_tensor(s);
  size_t n = acc.shape.num_elements();
  std::vector<float> h(n);
  for (size_t i = 0; i < n; ++i) h[i] = f(i);
  checkCUDA(cudaMemcpy(acc.ptr, h.data(), n * sizeof(float), cudaMemcpyHostToDevice));
  return acc;
}
static GenericTensorAccessorW random_filled(TensorShape const &s, Allocator &a) {
  std::uniform_real_distribution<float> d(-1.f, 1.f);
  return filled(s, a, [&](size_t) { return d(rng); });
}
static GenericTensorAccessorW const_filled(TensorShape const &s, Allocator &a, float v) {
  return filled(s, a, [&](size_t) { return v; });
}
static GenericTensorAccessorW iota_filled(TensorShape const &s, Allocator &a) {
  return filled(s, a, [&](size_t i) { return (float)i; });
}
static bool contains_non_zero(std::vector<float> const &v) {
  for (float x : v) if (x != 0.f) return true;
  return false;
}
static GenericTensorAccessorR ro(GenericTensorAccessorW const &w) { return read_only_accessor_from_write_accessor(w); }

int main() {
  ManagedFFStream managed_stream{};
  ManagedPerDeviceFFHandle managed_handle{size_t(8) << 20};
  Allocator allocator = create_local_cuda_memory_allocator();
  ffStream_t stream = managed_stream.raw_stream();
  PerDeviceFFHandle handle = managed_handle.raw_handle();
  auto sync = [&] { checkCUDA(cudaStreamSynchronize(stream)); };

  {  // test_replicate_kernel.cc:9-34 — forward must equal the input exactly
    TensorShape shape = make_float_tensor_shape_from_legion_dims({100});
    auto in = const_filled(shape, allocator, 1.0f);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Replicate::forward_kernel(stream, ro(in), out);
    sync();
    CHECK(to_host<float>(out.ptr, 100) == std::vector<float>(100, 1.0f));
    // backward: 10 stacked replicas of ones summed into a zero gradient
    auto og = const_filled(make_float_tensor_shape_from_legion_dims({100, 10}), allocator, 1.0f);
    auto ig = const_filled(shape, allocator, 0.0f);
    Kernels::Replicate::backward_kernel(stream, ig, ro(og), 10);
    sync();
    CHECK(to_host<float>(ig.ptr, 100) == std::vector<float>(100, 10.0f));
  }
  {  // test_reduction_kernel.cc:9-55 — 5 replicas of 10^5... (10 x 5 here keeps the sum exact)
    size_t n = 10, reps = 5;
    auto in = random_filled(make_float_tensor_shape_from_legion_dims({n, reps}), allocator);
    auto out = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({n}));
    Kernels::Reduction::forward_kernel(stream, ro(in), out, reps);
    sync();
    CHECK(contains_non_zero(to_host<float>(out.ptr, n)));
    auto og = const_filled(make_float_tensor_shape_from_legion_dims({n}), allocator, 1.0f);
    auto ig = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({n}));
    Kernels::Reduction::backward_kernel(stream, ig, ro(og));
    sync();
    CHECK(to_host<float>(ig.ptr, n) == std::vector<float>(n, 1.0f));  // pinned by the reference
  }
  {  // test_partition_kernel.cc:9-58 — forward copies 1s; backward 2 + 1 = 3
    TensorShape shape = make_float_tensor_shape_from_legion_dims({10, 10});
    auto st = Kernels::Repartition::init_kernel(handle, DataType::FLOAT);
    auto in = const_filled(shape, allocator, 1.0f);
    auto out = const_filled(shape, allocator, 0.0f);
    Kernels::Repartition::forward_kernel(stream, st, ro(in), out);
    sync();
    CHECK(to_host<float>(out.ptr, 100) == std::vector<float>(100, 1.0f));
    auto og = const_filled(shape, allocator, 1.0f);
    auto ig = const_filled(shape, allocator, 2.0f);
    Kernels::Repartition::backward_kernel(stream, st, ig, ro(og));
    sync();
    CHECK(to_host<float>(ig.ptr, 100) == std::vector<float>(100, 3.0f));
  }
  {  // test_combine_kernel.cc — 100 x 100
    TensorShape shape = make_float_tensor_shape_from_legion_dims({100, 100});
    auto in = random_filled(shape, allocator);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Combine::forward_kernel(stream, ro(in), out);
    sync();
    CHECK(to_host<float>(out.ptr, 10000) == to_host<float>(in.ptr, 10000));
    auto ig = const_filled(shape, allocator, 0.0f);
    Kernels::Combine::backward_kernel(stream, ro(out), ig);
    sync();
    CHECK(to_host<float>(ig.ptr, 10000) == to_host<float>(in.ptr, 10000));
  }
  {  // test_flat_kernel.cc:9-54 — forward 2s, backward 0 + 1 ... exact
    TensorShape shape = make_float_tensor_shape_from_legion_dims({100});
    auto in = const_filled(shape, allocator, 2.0f);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Flat::forward_kernel(stream, ro(in), out.get_float_ptr());
    sync();
    CHECK(to_host<float>(out.ptr, 100) == std::vector<float>(100, 2.0f));
    auto og = const_filled(shape, allocator, 0.0f);
    auto ig = const_filled(shape, allocator, 1.0f);
    Kernels::Flat::backward_kernel(stream, ro(in), ig.get_float_ptr(), og.get_float_ptr());
    sync();
    CHECK(to_host<float>(ig.ptr, 100) == std::vector<float>(100, 1.0f));
  }
  {  // test_reshape_kernel.cc:9-56
    TensorShape shape = make_float_tensor_shape_from_legion_dims({100});
    auto st = Kernels::Reshape::init_kernel(DataType::FLOAT);
    auto in = const_filled(shape, allocator, 1.0f);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Reshape::forward_kernel(stream, st, ro(in), out);
    sync();
    CHECK(to_host<float>(out.ptr, 100) == std::vector<float>(100, 1.0f));
    auto og = const_filled(shape, allocator, 1.0f);
    auto ig = const_filled(shape, allocator, 2.0f);
    Kernels::Reshape::backward_kernel(stream, st, ig, ro(og));
    sync();
    CHECK(to_host<float>(ig.ptr, 100) == std::vector<float>(100, 3.0f));
  }
  {  // test_softmax_kernel.cc — (n=1, c=100, h=1, w=1); backward is a copy (pinned)
    auto st = Kernels::Softmax::init_kernel(handle, 0, 1, 100, 1, 1);
    TensorShape shape = make_float_tensor_shape_from_legion_dims({100});
    auto in = random_filled(shape, allocator);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Softmax::forward_kernel(stream, st, in.get_float_ptr(), out.get_float_ptr());
    sync();
    auto h = to_host<float>(out.ptr, 100);
    float sum = 0;
    for (float x : h) sum += x;
    CHECK(contains_non_zero(h) && std::fabs(sum - 1.f) < 1e-5f);
    auto og = const_filled(shape, allocator, 1.0f);
    auto ig = allocator.allocate_tensor(shape);
    Kernels::Softmax::backward_kernel(stream, ig.get_float_ptr(), og.get_float_ptr(), 100);
    sync();
    CHECK(to_host<float>(ig.ptr, 100) == std::vector<float>(100, 1.0f));
  }
  {  // test_attention_kernel.cc:8-102
    size_t N = 10, H = 4, S = 64, L = 20;
    auto st = Kernels::MultiHeadAttention::init_kernel(handle, allocator, N, H, S, S, S, S, S, S, S, L, L, false);
    CHECK(st.weightSize == 4 * (3 * S * H * S + H * S * S));
    auto q = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto k = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto v = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto w = random_filled(make_float_tensor_shape_from_legion_dims({st.weightSize / 4}), allocator);
    auto out = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({L, N, S}));
    Kernels::MultiHeadAttention::forward_kernel(stream, st, q.get_float_ptr(), k.get_float_ptr(),
                                                v.get_float_ptr(), w.get_float_ptr(), out.get_float_ptr());
    sync();
    CHECK(contains_non_zero(to_host<float>(out.ptr, L * N * S)));
    auto qg = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto kg = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto vg = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    auto wg = const_filled(make_float_tensor_shape_from_legion_dims({st.weightSize / 4}), allocator, 0.f);
    auto og = random_filled(make_float_tensor_shape_from_legion_dims({L, N, S}), allocator);
    Kernels::MultiHeadAttention::backward_kernel(stream, st, q.get_float_ptr(), qg.get_float_ptr(),
                                                 k.get_float_ptr(), kg.get_float_ptr(), v.get_float_ptr(),
                                                 vg.get_float_ptr(), w.get_float_ptr(), wg.get_float_ptr(),
                                                 og.get_float_ptr());
    sync();
    CHECK(contains_non_zero(to_host<float>(wg.ptr, st.weightSize / 4)));
    Kernels::MultiHeadAttention::cleanup_kernel(allocator, st);
  }
  {  // test_batch_matmul_kernel.cc:8-75 — m = n = k = 10, batch 5
    size_t m = 10, n = 10, k = 10, b = 5;
    auto A = random_filled(make_float_tensor_shape_from_legion_dims({k, n, b}), allocator);
    auto B = random_filled(make_float_tensor_shape_from_legion_dims({m, k, b}), allocator);
    auto O = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({m, n, b}));
    Kernels::BatchMatmul::forward_kernel(stream, handle, O.get_float_ptr(), A.get_float_ptr(),
                                         B.get_float_ptr(), m, n, k, b, -1, -1, -1);
    sync();
    auto ho = to_host<float>(O.ptr, m * n * b), ha = to_host<float>(A.ptr, k * n * b), hb = to_host<float>(B.ptr, m * k * b);
    double ref = 0;
    for (size_t kk = 0; kk < k; ++kk) ref += (double)ha[kk] * hb[kk * m];
    CHECK(contains_non_zero(ho) && std::fabs(ho[0] - ref) < 1e-4);
    auto Og = random_filled(make_float_tensor_shape_from_legion_dims({m, n, b}), allocator);
    auto Ag = const_filled(make_float_tensor_shape_from_legion_dims({k, n, b}), allocator, 0.f);
    auto Bg = const_filled(make_float_tensor_shape_from_legion_dims({m, k, b}), allocator, 0.f);
    Kernels::BatchMatmul::backward_kernel(stream, handle, O.get_float_ptr(), Og.get_float_ptr(),
                                          A.get_float_ptr(), Ag.get_float_ptr(), B.get_float_ptr(),
                                          Bg.get_float_ptr(), m, n, k, b);
    sync();
    CHECK(contains_non_zero(to_host<float>(Ag.ptr, k * n * b)) && contains_non_zero(to_host<float>(Bg.ptr, m * k * b)));
  }
  {  // test_layer_norm_kernels.cc:8-73 — 10 x 10, eps 1e-5, affine
    size_t batch = 10, feat = 10;
    TensorShape shape = make_float_tensor_shape_from_legion_dims({feat, batch});
    TensorShape fshape = make_float_tensor_shape_from_legion_dims({feat});
    auto st = Kernels::LayerNorm::init_kernel(handle, allocator, true, batch, feat, 1e-5f);
    auto in = random_filled(shape, allocator);
    auto gamma = const_filled(fshape, allocator, 1.0f);
    auto beta = const_filled(fshape, allocator, 0.0f);
    auto out = allocator.allocate_tensor(shape);
    Kernels::LayerNorm::forward_kernel(stream, st, ro(in), out, gamma, beta);
    sync();
    auto h = to_host<float>(out.ptr, 100);
    float rowsum = 0;
    for (size_t j = 0; j < feat; ++j) rowsum += h[j];
    CHECK(contains_non_zero(h) && std::fabs(rowsum) < 1e-4f);
    auto og = random_filled(shape, allocator);
    auto ig = const_filled(shape, allocator, 0.f);
    auto gg = allocator.allocate_tensor(fshape);
    auto bg = allocator.allocate_tensor(fshape);
    Kernels::LayerNorm::backward_kernel(stream, st, ro(og), ro(in), ig, ro(gamma), gg, bg);
    sync();
    CHECK(contains_non_zero(to_host<float>(ig.ptr, 100)) && contains_non_zero(to_host<float>(gg.ptr, feat)));
  }
  {  // test_batch_norm_kernel.cc — 1 x 10 x 10 x 10, relu
    auto st = Kernels::BatchNorm::init_kernel(handle, allocator, nullptr, 1, 10, 10, 10, true);
    TensorShape shape = make_float_tensor_shape_from_legion_dims({10, 10, 10, 1});
    TensorShape cshape = make_float_tensor_shape_from_legion_dims({10});
    auto in = random_filled(shape, allocator);
    auto scale = const_filled(cshape, allocator, 1.0f);
    auto bias = const_filled(cshape, allocator, 0.0f);
    auto out = allocator.allocate_tensor(shape);
    Kernels::BatchNorm::forward_kernel(stream, st, in.get_float_ptr(), out.get_float_ptr(), scale.get_float_ptr(), bias.get_float_ptr());
    sync();
    CHECK(contains_non_zero(to_host<float>(out.ptr, 1000)));
    auto og = random_filled(shape, allocator);
    auto ig = const_filled(shape, allocator, 0.f);
    auto sg = const_filled(cshape, allocator, 0.f);
    auto bgr = const_filled(cshape, allocator, 0.f);
    Kernels::BatchNorm::backward_kernel(stream, st, in.get_float_ptr(), og.get_float_ptr(), out.get_float_ptr(),
                                        ig.get_float_ptr(), scale.get_float_ptr(), sg.get_float_ptr(), bgr.get_float_ptr(), 1000);
    sync();
    CHECK(contains_non_zero(to_host<float>(ig.ptr, 1000)) && contains_non_zero(to_host<float>(sg.ptr, 10)));
  }
  {  // test_dropout.cc — 10 x 10, rate 0.1, seed 12345
    TensorShape shape = make_float_tensor_shape_from_legion_dims({10, 10});
    auto st = Kernels::Dropout::init_kernel(handle, 0.1f, 12345ull, ArrayShape(shape), allocator);
    auto in = random_filled(shape, allocator);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Dropout::forward_kernel(stream, st, in.get_float_ptr(), out.get_float_ptr());
    sync();
    CHECK(contains_non_zero(to_host<float>(out.ptr, 100)));
    auto ig = allocator.allocate_tensor(shape);
    Kernels::Dropout::backward_kernel(stream, st, out.get_float_ptr(), ig.get_float_ptr());
    sync();
    CHECK(contains_non_zero(to_host<float>(ig.ptr, 100)));
  }
  {  // test_transpose_kernel.cc / test_reverse_kernels.cc / test_cast_kernel.cc / test_concat / test_split
    TensorShape shape = make_float_tensor_shape_from_legion_dims({10, 10});
    auto st = Kernels::Transpose::init_kernel(2, {ff_dim_t{1}, ff_dim_t{0}});
    auto in = iota_filled(shape, allocator);
    auto out = allocator.allocate_tensor(shape);
    Kernels::Transpose::forward_kernel(stream, st, ro(in), out);
    sync();
    auto h = to_host<float>(out.ptr, 100);
    CHECK(h[1] == 10.f && h[10] == 1.f && h[99] == 99.f);
    auto rev = allocator.allocate_tensor(shape);
    Kernels::Reverse::forward_kernel(stream, in.get_float_ptr(), rev.get_float_ptr(), 1, 10, 10, 100);
    sync();
    auto hr = to_host<float>(rev.ptr, 100);
    CHECK(hr[0] == 90.f && hr[99] == 9.f);
    auto casted = allocator.allocate_tensor(make_tensor_shape_from_legion_dims({10, 10}, DataType::INT32));
    Kernels::Cast::forward_kernel(stream, ro(in), casted, DataType::FLOAT, DataType::INT32);
    sync();
    auto hc = to_host<int32_t>(casted.ptr, 100);
    CHECK(hc[7] == 7 && hc[99] == 99);
    auto cat = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({20, 10}));
    Kernels::Concat::forward_kernel(stream, cat, {ro(in), ro(out)}, ff_dim_t{1});
    sync();
    auto hcat = to_host<float>(cat.ptr, 200);
    CHECK(hcat[0] == 0.f && hcat[9] == 9.f && hcat[10] == h[0] && hcat[20] == 10.f);
  }
  {  // element unary / binary / linear: no reference test exists; smoke through the API
    TensorShape shape = make_float_tensor_shape_from_legion_dims({64, 8});
    auto a = random_filled(shape, allocator), b = random_filled(shape, allocator);
    auto o = allocator.allocate_tensor(shape);
    auto eb = Kernels::ElementBinary::init_kernel(handle, OperatorType::EW_ADD, false, false, a.shape, b.shape, o.shape);
    Kernels::ElementBinary::forward_kernel(stream, eb, a.get_float_ptr(), b.get_float_ptr(), o.get_float_ptr(),
                                           OperatorType::EW_ADD, false, handle);
    sync();
    auto ha = to_host<float>(a.ptr, 512), hb = to_host<float>(b.ptr, 512), ho = to_host<float>(o.ptr, 512);
    bool ok = true;
    for (int i = 0; i < 512; ++i) ok &= ho[i] == ha[i] + hb[i];
    CHECK(ok);
    ElementUnaryAttrs attrs{OperatorType::RELU, std::nullopt};
    auto eu = Kernels::ElementUnary::init_kernel(a.shape, o.shape, attrs);
    Kernels::ElementUnary::forward_kernel(stream, eu, attrs, handle, ro(a), o);
    sync();
    ho = to_host<float>(o.ptr, 512);
    ok = true;
    for (int i = 0; i < 512; ++i) ok &= ho[i] == (ha[i] > 0 ? ha[i] : 0.f);
    CHECK(ok);
    // Linear 8 -> 16 over 64 rows, relu, L2
    auto W = random_filled(make_float_tensor_shape_from_legion_dims({8, 16}), allocator);
    auto bias = random_filled(make_float_tensor_shape_from_legion_dims({16}), allocator);
    auto Y = allocator.allocate_tensor(make_float_tensor_shape_from_legion_dims({16, 64}));
    auto lin = Kernels::Linear::init_kernel(handle, nullptr, Activation::RELU, RegularizerAttrs{L2RegularizerAttrs{1e-4f}},
                                            true, DataType::FLOAT, DataType::FLOAT, DataType::FLOAT, 64, 16);
    Kernels::Linear::forward_kernel(stream, lin, a.ptr, Y.ptr, W.ptr, bias.ptr, 8, 16, 64);
    sync();
    auto hw = to_host<float>(W.ptr, 128), hbias = to_host<float>(bias.ptr, 16), hy = to_host<float>(Y.ptr, 1024);
    double acc = hbias[3];
    for (int i = 0; i < 8; ++i) acc += (double)ha[5 * 8 + i] * hw[3 * 8 + i];
    CHECK(std::fabs(hy[5 * 16 + 3] - (acc > 0 ? acc : 0)) < 1e-5);
    auto dY = random_filled(make_float_tensor_shape_from_legion_dims({16, 64}), allocator);
    auto dX = const_filled(shape, allocator, 0.f);
    auto dW = const_filled(make_float_tensor_shape_from_legion_dims({8, 16}), allocator, 0.f);
    auto dB = const_filled(make_float_tensor_shape_from_legion_dims({16}), allocator, 0.f);
    Kernels::Linear::backward_kernel(stream, lin, a.ptr, dX.ptr, Y.ptr, dY.ptr, W.ptr, dW.ptr, dB.ptr, 8, 16, 64);
    sync();
    CHECK(contains_non_zero(to_host<float>(dW.ptr, 128)) && contains_non_zero(to_host<float>(dX.ptr, 512)));
  }
  {  // profiling_wrapper returns the TOTAL of the measured iterations; arena recycles blocks
    TensorShape shape = make_float_tensor_shape_from_legion_dims({1 << 20});
    auto in = const_filled(shape, allocator, 1.0f);
    auto out = allocator.allocate_tensor(shape);
    auto f = [&](ffStream_t s, GenericTensorAccessorR const &i, GenericTensorAccessorW const &o) {
      Kernels::Combine::forward_kernel(s, i, o);
    };
    auto t1 = profiling_wrapper(f, ProfilingSettings{2, 1}, ro(in), out);
    auto t8 = profiling_wrapper(f, ProfilingSettings{2, 8}, ro(in), out);
    CHECK(t1.has_value() && t8.has_value() && *t1 > 0.f && *t8 > *t1);
    auto none = profiling_wrapper(f, false, ro(in), out);
    CHECK(!none.has_value());
    LocalCudaAllocator arena;
    void *p1 = arena.allocate(1000);
    arena.deallocate(p1);
    void *p2 = arena.allocate(900);
    CHECK(p1 == p2 && arena.driver_allocations() == 1 && arena.bytes_in_use() == 1024);
    bool threw = false;
    try { (void)in.get<DataType::INT32>(); } catch (std::runtime_error const &) { threw = true; }
    CHECK(threw);  // accessor dtype mismatch throws, like the reference (accessor.h:16-23)
  }
  std::printf("%d checks, %d failed\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
