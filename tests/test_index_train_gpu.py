"""GPU parity: Embedding / Gather / TopK / losses / optimizers / metrics vs the oracle."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_IDX, SEED_W, assert_close, assert_equal_bits, to_dev, uniform

pytestmark = pytest.mark.gpu


def _indices(shape, hi, seed=SEED_IDX, zipf=False):
    rng = np.random.default_rng(seed)
    if zipf:
        return np.minimum(rng.zipf(1.05, size=shape) - 1, hi - 1).astype(np.int64)
    return rng.integers(0, hi, size=shape, dtype=np.int64)


@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("entries,out_dim,batch", [(1000, 256, 777), (50, 64, 1), (333, 8, 4096)])
def test_embedding_no_aggr_rows_are_bit_exact(K, oracle, idx_dtype, dtype, entries, out_dim, batch):
    table, tn = to_dev(uniform((entries, out_dim), SEED_W), dtype)
    idx_np = _indices((batch,), entries)
    idx = torch.from_numpy(idx_np).cuda().to(idx_dtype)
    out = torch.zeros(batch, out_dim, dtype=dtype, device="cuda")
    K.Embedding.forward_kernel(None, idx, out, table, K.AGGR_NONE, 1, out_dim, batch)
    assert_equal_bits(out, table[idx.long()], "embedding gather")   # bit-exact index path
    ref = np.zeros((batch, out_dim), np.float32)
    oracle.embedding_forward(idx_np, ref, tn, -1, 1, out_dim, batch)
    assert (out.float().cpu().numpy() == ref).all()
    dout, dn = to_dev(uniform((batch, out_dim), SEED_ACT), dtype)
    dtab, dtn = to_dev(uniform((entries, out_dim), SEED_W + 1), dtype)
    K.Embedding.backward_kernel(None, idx, dout, dtab, K.AGGR_NONE, 1, out_dim, batch)
    oracle.embedding_backward(idx_np, dn, dtn, -1, 1, out_dim, batch)
    # 16-bit atomics round after every add; rows hit ~batch/entries times
    rt = None if dtype == torch.float32 else 2e-2 * max(1.0, np.sqrt(batch / entries))
    assert_close(dtab, dtn, dtype, rtol=rt, what="embedding bwd")


@pytest.mark.parametrize("aggr", ["SUM", "AVG"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_embedding_bags(K, oracle, aggr, dtype):
    entries, out_dim, batch, bag = 500, 64, 300, 7
    code = getattr(K, "AGGR_" + aggr)
    table, tn = to_dev(uniform((entries, out_dim), SEED_W), dtype)
    idx_np = _indices((batch, bag), entries)
    idx = torch.from_numpy(idx_np).cuda()
    out = torch.zeros(batch, out_dim, dtype=dtype, device="cuda")
    K.Embedding.forward_kernel(None, idx, out, table, code, bag, out_dim, batch)
    ref = np.zeros((batch, out_dim), np.float32)
    oracle.embedding_forward(idx_np, ref, tn, code, bag, out_dim, batch)
    assert_close(out, ref, dtype, what=f"embedding {aggr}")
    dout, dn = to_dev(uniform((batch, out_dim), SEED_ACT), dtype)
    dtab, dtn = to_dev(np.zeros((entries, out_dim), np.float32), dtype)
    K.Embedding.backward_kernel(None, idx, dout, dtab, code, bag, out_dim, batch)
    oracle.embedding_backward(idx_np, dn, dtn, code, bag, out_dim, batch)
    assert_close(dtab, dtn, dtype, rtol=None if dtype == torch.float32 else 6e-2, what=f"embedding bwd {aggr}")


def test_embedding_full_size_table(K):
    """config[1]: 1M x 256 table, 65536 int64 indices (uniform and Zipf). Forward rows must be
    bit-identical to the table rows; backward is checked by column sums (sum of all scattered
    gradient equals the sum of dOut, independent of the index distribution)."""
    entries, out_dim, batch = 1_000_000, 256, 65536
    g = torch.Generator(device="cuda").manual_seed(SEED_W)
    table = torch.rand(entries, out_dim, device="cuda", generator=g)
    for zipf in (False, True):
        idx = torch.from_numpy(_indices((batch,), entries, zipf=zipf)).cuda()
        out = torch.empty(batch, out_dim, device="cuda")
        K.Embedding.forward_kernel(None, idx, out, table, K.AGGR_NONE, 1, out_dim, batch)
        assert torch.equal(out, table[idx])
        dtab = torch.zeros(entries, out_dim, device="cuda")
        K.Embedding.backward_kernel(None, idx, out, dtab, K.AGGR_NONE, 1, out_dim, batch)
        got, want = dtab.double().sum(0), out.double().sum(0)
        # fp32 atomics accumulate sequentially: a Zipf-hot row receives ~1e4 addends
        assert ((got - want).abs() / want.abs()).max().item() < (2e-3 if zipf else 1e-5)
        touched = torch.zeros(entries, dtype=torch.bool, device="cuda")
        touched[idx] = True
        assert dtab[~touched].abs().max().item() == 0


def test_rand_generate_int(K):
    for dt in (torch.int32, torch.int64):
        t = torch.zeros(1000, dtype=dt, device="cuda")
        K.Embedding.rand_generate_int(None, t, 37)
        assert torch.equal(t, (torch.arange(1000, device="cuda") % 37).to(dt))


@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
@pytest.mark.parametrize("shape,dim,odim", [((4, 6, 5), 1, 3), ((10, 7), 0, 12), ((3, 4, 5, 6), 3, 2)])
def test_gather(K, oracle, idx_dtype, shape, dim, odim):
    x, xn = to_dev(uniform(shape, SEED_ACT))
    oshape = list(shape)
    oshape[dim] = odim
    idx_np = _indices(tuple(oshape), shape[dim])
    idx = torch.from_numpy(idx_np).cuda().to(idx_dtype)
    out = torch.zeros(oshape, device="cuda")
    K.Gather.forward_kernel(None, dim, x, idx, out)
    assert torch.equal(out, torch.gather(x, dim, idx.long()))
    n, stride, idim, od = K.Gather._geom(x, out, dim)
    ref = np.zeros(oshape, np.float32)
    oracle.gather_forward(xn, idx_np, ref, n, stride, idim, od)
    assert (out.cpu().numpy() == ref).all()
    dout, dn = to_dev(uniform(tuple(oshape), SEED_ACT + 1))
    din, dinn = to_dev(uniform(shape, SEED_ACT + 2))
    K.Gather.backward_kernel(None, dim, dout, idx, din)
    oracle.gather_backward(dn, idx_np, dinn, n, stride, idim, od)
    assert_close(din, dinn, what="gather bwd")


@pytest.mark.parametrize("batch,length,k", [(5, 100, 10), (3, 1000, 1), (2, 64, 64), (7, 5000, 17)])
def test_topk(K, oracle, batch, length, k):
    xin = uniform((batch, length), SEED_ACT)
    xin[:, ::7] = 0.5  # ties: the lower index must come first
    x, xn = to_dev(xin)
    out = torch.zeros(batch, k, device="cuda")
    ind = torch.zeros(batch, k, dtype=torch.int32, device="cuda")
    K.TopK.forward_kernel(None, x, out, ind, batch, length, k)
    ro, ri = np.zeros((batch, k), np.float32), np.zeros((batch, k), np.int32)
    oracle.topk_forward(xn, ro, ri, batch, length, k)
    assert (out.cpu().numpy() == ro).all() and (ind.cpu().numpy() == ri).all()
    dout, dn = to_dev(uniform((batch, k), SEED_ACT + 1))
    din, dinn = to_dev(uniform((batch, length), SEED_ACT + 2))
    K.TopK.backward_kernel(None, dout, ind, din, batch, length, k)
    oracle.topk_backward(dn, ri, dinn, batch, length, k)
    assert_close(din, dinn, what="topk bwd")


@pytest.mark.parametrize("k", [1, 2])
def test_sparse_cce_loss(K, oracle, k):
    rows, classes = 64, 10  # config[0]: 64 x 10 logits, int32 labels
    p, pn = to_dev(uniform((rows, classes), SEED_ACT, 0, 1))
    lab_np = np.random.default_rng(SEED_IDX).integers(0, classes, rows // k).astype(np.int32)
    lab = torch.from_numpy(lab_np).cuda()
    g = torch.zeros_like(p)
    scale = 1.0 / rows
    K.sparse_categorical_crossentropy_loss_backward_kernel(None, g, p, lab, rows, classes, k, scale)
    ref = np.zeros_like(pn)
    oracle.sparse_cce_loss_backward(ref, pn, lab_np, rows, classes, k, scale)
    assert_close(g, ref, rtol=1e-6, what="sparse cce")


def test_dense_losses(K, oracle):
    n = 5000
    p, pn = to_dev(uniform((n,), SEED_ACT))
    y, yn = to_dev(uniform((n,), SEED_W))
    ref = np.zeros(n, np.float32)
    oracle.diff_loss_backward(ref, pn, yn, n, 0.125)
    for fn in (K.categorical_crossentropy_loss_backward_kernel, K.mean_squared_error_avg_loss_backward_kernel):
        g = torch.zeros_like(p)
        fn(None, g, p, y, 0.125)
        assert (g.cpu().numpy() == ref).all()
    g = torch.zeros_like(p)
    K.identity_loss_backward_kernel(None, g, p, 0.25)
    assert (g == 0.25).all()


@pytest.mark.parametrize("momentum,nesterov", [(0.0, False), (0.9, False), (0.9, True)])
@pytest.mark.parametrize("reps", [1, 3])
def test_sgd(K, oracle, momentum, nesterov, reps):
    n = 669_706  # parameter count of the config[0] MLP
    w, wn = to_dev(uniform((n,), SEED_W))
    g, gn = to_dev(uniform((reps, n), SEED_ACT))
    v, vn = to_dev(uniform((n,), SEED_W + 1) * 0.1)
    for _ in range(2):
        K.sgd_ps_update_task_gpu(None, 0.01, momentum, nesterov, 1e-4, g, n, reps, w, v)
        oracle.sgd_update(0.01, momentum, int(nesterov), 1e-4, gn, n, reps, wn, vn)
    assert_close(w, wn, rtol=1e-6, what="sgd W")
    if momentum > 0:
        assert_close(v, vn, rtol=1e-6, what="sgd V")


def test_adam_and_master_weights(K, oracle):
    n = 100_003
    w, wn = to_dev(uniform((n,), SEED_W))
    g, gn = to_dev(uniform((2, n), SEED_ACT))
    m, mn = to_dev(np.zeros(n, np.float32))
    v, vn = to_dev(np.zeros(n, np.float32))
    for t in range(1, 4):
        alpha_t = 1e-3 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        K.adam_ps_update_task_gpu(None, alpha_t, 0.9, 0.999, 1e-4, 1e-8, g, n, 2, m, v, w)
        oracle.adam_update(alpha_t, 0.9, 0.999, 1e-4, 1e-8, gn, n, 2, mn, vn, wn)
    assert_close(w, wn, rtol=1e-5, what="adam W")
    assert_close(m, mn, rtol=1e-5, what="adam M")
    assert_close(v, vn, rtol=1e-5, what="adam V")
    # bf16 weights with an fp32 master copy
    master, master_n = to_dev(uniform((n,), SEED_W))
    wb = master.to(torch.bfloat16)
    gb, gbn = to_dev(uniform((1, n), SEED_ACT), torch.bfloat16)
    K.sgd_ps_update_task_gpu(None, 0.1, 0.0, False, 0.0, gb, n, 1, wb, None, master)
    oracle.sgd_update(0.1, 0.0, 0, 0.0, gbn, n, 1, master_n, None)
    assert_close(master, master_n, rtol=1e-6, what="master weights")
    assert torch.equal(wb, master.to(torch.bfloat16))


def test_metrics(K, oracle):
    rows, classes = 1000, 10
    rng = np.random.default_rng(SEED_ACT)
    logits = rng.random((rows, classes), dtype=np.float32)
    logits /= logits.sum(1, keepdims=True)
    lab = rng.integers(0, classes, rows).astype(np.int32)
    flags = K.METRIC_ACCURACY | K.METRIC_SPARSE_CCE | K.METRIC_MSE | K.METRIC_RMSE | K.METRIC_MAE
    perf = K.PerfMetrics()
    lt, lb = torch.from_numpy(logits).cuda(), torch.from_numpy(lab).cuda()
    K.update_metrics_sparse_label_kernel(None, flags, lt, lb, rows, classes, perf)
    K.update_metrics_sparse_label_kernel(None, flags, lt, lb, rows, classes, perf)  # accumulates
    ref = np.zeros(7, np.float64)
    oracle.metrics_sparse(flags, logits, lab, rows, classes, ref)
    got = perf.read()
    assert got["train_all"] == 2 * int(ref[0]) and got["train_correct"] == 2 * int(ref[1])
    for key, i in (("sparse_cce_loss", 3), ("mse_loss", 4), ("rmse_loss", 5), ("mae_loss", 6)):
        assert abs(got[key] - 2 * ref[i]) <= 1e-4 * abs(2 * ref[i]), key
    onehot = np.eye(classes, dtype=np.float32)[lab]
    flags = K.METRIC_ACCURACY | K.METRIC_CCE | K.METRIC_MSE | K.METRIC_RMSE | K.METRIC_MAE
    perf = K.PerfMetrics()
    K.update_metrics_label_kernel(None, flags, lt, torch.from_numpy(onehot).cuda(), rows, classes, perf)
    ref = np.zeros(7, np.float64)
    oracle.metrics_label(flags, logits, onehot, rows, classes, ref)
    got = perf.read()
    assert got["train_all"] == int(ref[0]) and got["train_correct"] == int(ref[1])
    for key, i in (("cce_loss", 2), ("mse_loss", 4), ("rmse_loss", 5), ("mae_loss", 6)):
        assert abs(got[key] - ref[i]) <= 1e-4 * abs(ref[i]), key
