// The cost-estimator bridge (flexflow-train_b200/include/ffk_cpp/cost_estimator.h) used the way
// the reference's machine-mapping search uses a CostEstimator
// (lib/compiler/src/compiler/machine_mapping/get_optimal_machine_mapping.cc:159,243):
// estimate_cost(OpCostEstimateKey) for every operator family, estimate_cost(TensorSetMovement) for
// the edges. Built by __graft_entry__.build(), run on the GPU by tests/test_cpp_api_gpu.py.
#include <cstdio>

#include "ffk_cpp/cost_estimator.h"

using namespace FlexFlow;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                       \
  do {                                                                    \
    ++g_checks;                                                           \
    if (!(cond)) {                                                        \
      ++g_fail;                                                           \
      std::printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
    }                                                                     \
  } while (0)

static OpCostEstimateKey key(PCGOperatorAttrs a, std::vector<ParallelTensorShape> in,
                             std::vector<ParallelTensorShape> w, std::vector<ParallelTensorShape> out) {
  return OpCostEstimateKey{std::move(a), std::move(in), std::move(w), std::move(out), MachineView{0, {1}}};
}

int main() {
  CostEstimator est = get_local_cost_estimator(ProfilingSettings{1, 3});
  LocalCostEstimator local(ProfilingSettings{1, 3});
  auto P = [](std::vector<size_t> d, DataType dt = DataType::FLOAT) { return lift_to_parallel(d, dt); };
  struct Case {
    char const *name;
    OpCostEstimateKey k;
  };
  std::vector<Case> cases;
  cases.push_back({"linear", key(LinearAttrs{512, true, DataType::FLOAT, Activation::RELU, std::nullopt},
                                 {P({256, 384})}, {P({384, 512}), P({512})}, {P({256, 512})})});
  cases.push_back({"batch_matmul", key(BatchMatmulAttrs{}, {P({8, 64, 32}), P({8, 32, 48})}, {}, {P({8, 64, 48})})});
  cases.push_back({"multi_head_attention", key(MultiHeadAttentionAttrs{64, 4, 16, 16}, {P({4, 20, 64}), P({4, 20, 64}), P({4, 20, 64})},
                                               {}, {P({4, 20, 64})})});
  cases.push_back({"element_unary", key(ElementUnaryAttrs{OperatorType::GELU, std::nullopt}, {P({128, 256})}, {}, {P({128, 256})})});
  cases.push_back({"element_binary", key(ElementBinaryAttrs{OperatorType::EW_ADD}, {P({128, 256}), P({128, 256})}, {}, {P({128, 256})})});
  cases.push_back({"softmax", key(SoftmaxAttrs{ff_dim_t{1}}, {P({128, 256})}, {}, {P({128, 256})})});
  cases.push_back({"layer_norm", key(LayerNormAttrs{{ff_dim_t{1}}, true, 1e-5f}, {P({128, 256})}, {}, {P({128, 256})})});
  cases.push_back({"batch_norm", key(BatchNormAttrs{true}, {P({4, 8, 12, 12})}, {}, {P({4, 8, 12, 12})})});
  cases.push_back({"dropout", key(DropoutAttrs{0.1f, 12345ull}, {P({128, 256})}, {}, {P({128, 256})})});
  cases.push_back({"embedding", key(EmbeddingAttrs{1000, 64, std::nullopt}, {P({512}, DataType::INT64)}, {P({1000, 64})}, {P({512, 64})})});
  cases.push_back({"reduce", key(ReduceAttrs{{ff_dim_t{1}}, OperatorType::REDUCE_SUM}, {P({128, 256})}, {}, {P({128, 1})})});
  cases.push_back({"pool2d", key(Pool2DAttrs{2, 2, 2, 2, 0, 0, PoolOp::MAX, std::nullopt}, {P({4, 8, 12, 12})}, {}, {P({4, 8, 6, 6})})});
  cases.push_back({"conv2d", key(Conv2DAttrs{16, 3, 3, 1, 1, 1, 1, 1, std::nullopt, true}, {P({4, 8, 12, 12})},
                                 {P({16, 8, 3, 3}), P({16})}, {P({4, 16, 12, 12})})});
  cases.push_back({"cast", key(CastAttrs{DataType::HALF}, {P({128, 256})}, {}, {P({128, 256}, DataType::HALF)})});
  cases.push_back({"transpose", key(TransposeAttrs{{ff_dim_t{1}, ff_dim_t{0}}}, {P({128, 256})}, {}, {P({256, 128})})});
  for (Case const &c : cases) {
    float ms = est.estimate_cost(c.k);
    CostDetails d = local.estimate_cost(c.k.op_attrs, c.k.input_shapes, c.k.weight_shapes, c.k.output_shapes, c.k.machine_view);
    std::printf("%-22s %8.4f ms  %9zu bytes\n", c.name, ms, d.total_mem_usage);
    CHECK(ms > 0.f && ms < 1000.f);
    CHECK(d.total_mem_usage > 0);
  }
  // a bigger Linear costs more than a smaller one, and a tensor-parallel PIECE costs less than the whole
  auto lin = [&](size_t batch, int out, int degree) {
    ParallelTensorShape o = P({batch, size_t(out)}), w = P({1024, size_t(out)});
    o.shard_dims[1].degree = degree;
    w.shard_dims[1].degree = degree;
    return est.estimate_cost(key(LinearAttrs{out, false, DataType::FLOAT, std::nullopt, std::nullopt}, {P({batch, 1024})}, {w}, {o}));
  };
  float whole = lin(4096, 4096, 1), half = lin(4096, 4096, 2), tiny = lin(64, 128, 1);
  std::printf("linear 4096x1024x4096: whole %.3f ms, out-channels / 2 %.3f ms, 64x1024x128 %.3f ms\n", whole, half, tiny);
  CHECK(whole > tiny);
  CHECK(half < whole);
  // the free operators (local_cost_estimator.cc:36-39)
  CHECK(est.estimate_cost(key(ReplicateAttrs{8}, {P({64, 64})}, {}, {P({64, 64})})) == 0.f);
  CHECK(est.estimate_cost(key(RepartitionAttrs{ff_dim_t{0}, 8}, {P({64, 64})}, {}, {P({64, 64})})) == 0.f);
  CHECK(est.estimate_cost(key(InputAttrs{}, {}, {}, {P({64, 64})})) == 0.f);
  CHECK(est.estimate_cost(key(WeightAttrs{}, {}, {}, {P({64, 64})})) == 0.f);
  // movement: nothing for identical views, a copy for one destination, growing with the fan-out
  MachineView a{0, {1}}, b{1, {1}}, fan{0, {8}};
  SingleTensorMovement same{P({4096, 4096}), {a}, {a}}, copy{P({4096, 4096}), {a}, {b}}, spread{P({4096, 4096}), {a}, {fan}};
  float t_same = est.estimate_cost(TensorSetMovement{{same}}), t_copy = est.estimate_cost(TensorSetMovement{{copy}}),
        t_fan = est.estimate_cost(TensorSetMovement{{spread}}), t_both = est.estimate_cost(TensorSetMovement{{copy, spread}});
  std::printf("movement of 64 MiB: same view %.4f ms, one peer %.4f ms, 8 devices %.4f ms\n", t_same, t_copy, t_fan);
  CHECK(t_same == 0.f);
  CHECK(t_copy > 0.08f && t_copy < 0.12f);   // 67 MB at 770 GB/s + launch
  CHECK(t_fan > t_copy);
  CHECK(t_both == t_fan);                    // concurrent movements: the slowest decides
  std::printf("%d checks, %d failed\n", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
