"""GPU parity of whole training steps through the step driver (flexflow-train_b200/training.py):
BASELINE configs[0] — MLP 784-512-512-10, batch 64, fp32, softmax + sparse-CCE + SGD — against
the same chain of CPU-oracle kernels, step by step; and a small encoder stack for finiteness and
loss decrease."""
import numpy as np
import pytest
import torch

from helpers import assert_close, assert_close_l2, uniform

pytestmark = pytest.mark.gpu


def oracle_mlp_step(o, P, x, labels, lr, wd, batch):
    """The same step on the CPU oracle. P: list of (W, b) numpy arrays (updated in place)."""
    acts = [x]
    dims = [x.shape[1]] + [w.shape[0] for w, _ in P]
    for i, (w, b) in enumerate(P):
        y = np.zeros((batch, w.shape[0]), np.float32)
        o.linear_forward(acts[-1], w, b, y, w.shape[1], w.shape[0], batch, -1 if i == len(P) - 1 else 0)
        acts.append(y)
    probs = np.zeros_like(acts[-1])
    o.softmax_forward(acts[-1], probs, batch, probs.shape[1], 1)
    g = np.zeros_like(probs)
    o.sparse_cce_loss_backward(g, probs, labels, batch, probs.shape[1], 1, 1.0 / batch)
    # softmax backward = copy (reference semantics)
    grads = []
    for i in reversed(range(len(P))):
        w, b = P[i]
        dw, db = np.zeros_like(w), np.zeros_like(b)
        dx = np.zeros_like(acts[i])
        o.linear_backward(acts[i], dx, acts[i + 1], g, w, dw, db, w.shape[1], w.shape[0], batch,
                          -1 if i == len(P) - 1 else 0, 0.0, None)
        grads.append((dw, db))
        g = dx
    for (w, b), (dw, db) in zip(P, reversed(grads)):
        o.sgd_update(lr, 0.0, 0, wd, dw, w.size, 1, w.reshape(-1), None)
        o.sgd_update(lr, 0.0, 0, wd, db, b.size, 1, b, None)
    return probs


def test_mlp_config0_three_steps_match_the_oracle(ffk, oracle):
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    batch = 64
    model = T.build_mlp(batch=batch)
    opt = T.SGD(lr=0.01, weight_decay=1e-4)
    lin = [l for l in model.layers if isinstance(l, T.Linear)]
    P = [(l.w.cpu().numpy().copy(), l.b.cpu().numpy().copy()) for l in lin]
    rng = np.random.default_rng(1234)
    for step in range(3):
        x = rng.uniform(0, 1, (batch, 784)).astype(np.float32)       # MNIST-shaped synthetic data
        labels = rng.integers(0, 10, batch).astype(np.int32)
        model.input.data.copy_(torch.from_numpy(x))
        T.sparse_cce_step(model, torch.from_numpy(labels).cuda(), opt)
        probs = oracle_mlp_step(oracle, P, x, labels, 0.01, 1e-4, batch)
        assert_close(model.output.data, probs, what=f"step {step} probabilities")
        for l, (w, b) in zip(lin, P):
            assert_close(l.w, w, rtol=1e-4, what=f"step {step} weights")
            assert_close(l.b, b, rtol=1e-4, what=f"step {step} bias")


@pytest.mark.parametrize("fuse", [False, True])
@pytest.mark.parametrize("dtype,rate", [(torch.float32, 0.0), (torch.float32, 0.1), (torch.bfloat16, 0.1)])
def test_encoder_step_matches_the_oracle(ffk, oracle, dtype, rate, fuse):
    """BASELINE configs[2] at toy size: one full training step (forward, MSE gradient, backward with
    dq + dk + dv, SGD) of the transformer-encoder graph against the same step composed from the
    CPU oracle (tests/oracle_encoder.py), with the dropout masks the GPU drew. fp32: rtol 1e-4;
    bf16: the model output within 2e-2, updated weights within one bf16 rounding of the update."""
    import importlib
    from oracle_encoder import OracleEncoder
    T = importlib.import_module("flexflow_train_b200.training")
    B, S, D, H, hd, FF, L = 4, 16, 64, 4, 24, 128, 2
    model = T.build_transformer_encoder(num_layers=L, d_model=D, heads=H, seq=S, batch=B, dim_ff=FF, head_dim=hd,
                                        dropout=rate, dtype=dtype, fuse=fuse)
    per = len(model.layers) // L
    names = ("mha", "g1", "b1", "w1", "c1", "w2", "c2", "g2", "b2")

    def params(i):
        if fuse:   # fused-operator execution: the dropout states live inside the two dense layers
            att, ln1, ff1, ff2, ln2 = model.layers[i * per:(i + 1) * per]
            drops = (ff1.drop, ff2.drop)
        else:
            att, _, ln1, ff1, dp1, ff2, dp2, _, ln2 = model.layers[i * per:(i + 1) * per]
            drops = (dp1.state, dp2.state)
        return (att.w, ln1.gamma, ln1.beta, ff1.w, ff1.b, ff2.w, ff2.b, ln2.gamma, ln2.beta), drops

    g = torch.Generator(device="cuda").manual_seed(11)
    for i in range(L):   # non-trivial affine parameters and biases
        for t in params(i)[0][1:]:
            if t.dim() == 1:
                t.add_(((torch.rand(t.shape, device="cuda", generator=g) - 0.5) * 0.4).to(dtype))
    W = [{n: t.float().cpu().numpy().copy() for n, t in zip(names, params(i)[0])} for i in range(L)]
    x = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(dtype)
    target = (torch.rand(B, S, D, device="cuda", generator=g) * 2 - 1).to(dtype)
    opt = T.SGD(lr=0.05, weight_decay=1e-3)
    model.input.data.copy_(x)
    T.mse_step(model, target, opt)
    torch.cuda.synchronize()
    masks = [tuple(d.mask.cpu().numpy().view(np.uint32).copy() for d in params(i)[1]) for i in range(L)] if rate > 0 else None
    enc = OracleEncoder(oracle, W, B, S, D, H, hd, FF, rate=rate)
    out = enc.step(x.float().cpu().numpy(), target.float().cpu().numpy(), 0.05, 1e-3, masks=masks)
    if dtype == torch.float32:
        assert_close(model.output.data, out, what="encoder output")
    else:   # chains of 16-bit GEMMs: bounded in relative L2 (helpers.assert_close_l2)
        assert_close_l2(model.output.data, out, what="encoder output")
    for i in range(L):
        for n, t in zip(names, params(i)[0]):
            if dtype == torch.float32:
                assert_close(t, W[i][n], rtol=1e-4, what=f"layer {i} {n} after SGD")
            else:
                # master weights are fp32; the bf16 copy is within one rounding of the oracle's value
                assert_close(t, W[i][n], dtype, what=f"layer {i} {n} after SGD")


def test_small_encoder_trains(ffk):
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    for dtype in (torch.float32, torch.bfloat16):
        model = T.build_transformer_encoder(num_layers=2, d_model=64, heads=4, seq=16, batch=4, dim_ff=128,
                                            dropout=0.0, dtype=dtype)
        opt = T.SGD(lr=0.05, weight_decay=0.0)
        g = torch.Generator(device="cuda").manual_seed(1)
        x = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(dtype)
        target = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(dtype)
        losses = []
        for _ in range(8):
            model.input.data.copy_(x)
            T.mse_step(model, target, opt)
            losses.append(((model.output.data.float() - target.float()) ** 2).mean().item())
        assert all(np.isfinite(losses)), losses
        assert losses[-1] < losses[0], losses


def test_mlp_step_is_cuda_graph_capturable(ffk):
    """The kernel API must be capturable when profiling is off: capture a whole
    config[0] MLP training step (zero-grad, forward, loss gradient, backward, SGD) in a CUDA graph
    and check that replays advance the weights exactly like eager steps."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    batch = 64
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(batch, 784, device="cuda", generator=g)
    labels = torch.randint(0, 10, (batch,), device="cuda", generator=g, dtype=torch.int32)

    def run(use_graph):
        model = T.build_mlp(batch=batch)
        opt = T.SGD(lr=0.01, weight_decay=1e-4)
        model.input.data.copy_(x)
        T.sparse_cce_step(model, labels, opt)  # warm-up outside capture (lazy init, attributes)
        if use_graph:
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=s):
                    T.sparse_cce_step(model, labels, opt)
            torch.cuda.current_stream().wait_stream(s)
            for _ in range(3):
                graph.replay()
        else:
            for _ in range(3):
                T.sparse_cce_step(model, labels, opt)
        torch.cuda.synchronize()
        return [l.w.clone() for l in model.layers if isinstance(l, T.Linear)]

    eager, graphed = run(False), run(True)
    for a, b in zip(eager, graphed):
        assert torch.equal(a, b)


def test_direct_dropout_gradient_matches_the_accumulating_path(ffk):
    """A Dropout that is the only reader of its input writes dX straight into x.grad (no zero
    fill, no accumulate pass). Two steps must give bit-identical weights to the generic path."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")

    def run(direct):
        model = T.build_transformer_encoder(num_layers=2, d_model=64, heads=4, seq=16, batch=4, dim_ff=128,
                                            dropout=0.25, dtype=torch.bfloat16)
        n_direct = sum(1 for l in model.layers if isinstance(l, T.Dropout) and l.sole_reader)
        assert n_direct == 4
        if not direct:
            for l in model.layers:
                if isinstance(l, T.Dropout):
                    l.sole_reader = False
            model._overwritten.clear()
        opt = T.SGD(lr=0.05, weight_decay=0.0)
        g = torch.Generator(device="cuda").manual_seed(1)
        x = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
        target = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
        for _ in range(2):
            model.input.data.copy_(x)
            T.mse_step(model, target, opt)
        torch.cuda.synchronize()
        return [p.clone() for p in model.params]

    for a, b in zip(run(True), run(False)):
        assert torch.equal(a, b)


def test_small_bert_trains(ffk):
    """lib/models' BERT graph at toy size (GELU feed-forward, LayerNorm eps 1e-12, a vocabulary
    that is not a multiple of 8 so that the head's gradients take the repacked-operand path)."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    K = importlib.import_module("flexflow_train_b200.kernels")
    K.set_gemm_workspace(32 << 20)
    try:
        model = T.build_bert(num_layers=2, hidden=64, heads=4, dim_ff=128, seq=16, batch=4, vocab=203,
                             dropout=0.0)
        opt = T.SGD(lr=0.2, weight_decay=0.0)
        g = torch.Generator(device="cuda").manual_seed(1)
        x = (torch.rand(4, 16, 64, device="cuda", generator=g) * 2 - 1).to(torch.bfloat16)
        labels = torch.randint(0, 203, (4 * 16,), device="cuda", generator=g, dtype=torch.int32)
        losses = []
        for _ in range(10):
            model.input.data.copy_(x)
            T.sparse_cce_step(model, labels, opt)
            p = model.output.data.float().view(64, 203)
            losses.append(float(-torch.log(p[torch.arange(64), labels.long()].clamp_min(1e-8)).mean()))
        assert all(np.isfinite(losses)), losses
        assert losses[-1] < losses[0], losses
    finally:
        K.set_gemm_workspace(None)


def test_small_dlrm_trains(ffk):
    """examples/cpp/DLRM's graph at toy size on one GPU: loss must fall and only looked-up table
    rows may change."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    net = T.DLRM(batch=256, embedding_size=(1000,) * 4, sparse_feature_size=16, bag=2, mlp_bot=(4, 32, 16),
                 mlp_top=(32, 16, 2))
    opt = T.SGD(lr=0.5, weight_decay=0.0)
    g = torch.Generator(device="cuda").manual_seed(2)
    net.dense.data.copy_(torch.rand(256, 4, device="cuda", generator=g))
    target = (torch.rand(256, 2, device="cuda", generator=g) > 0.5).float()
    before = [t.clone() for t in net.emb.tables]
    losses = []
    for _ in range(30):
        T.dlrm_step(net, target, opt)
        losses.append(float(((net.model.output.data - target) ** 2).mean()))
    assert all(np.isfinite(losses)) and losses[-1] < losses[0], losses
    for j, (b, a) in enumerate(zip(before, net.emb.tables)):
        touched = torch.zeros(1000, dtype=torch.bool, device="cuda")
        touched[net.emb.idx[j].reshape(-1)] = True
        assert torch.equal(a[~touched], b[~touched]) and not torch.equal(a[touched], b[touched])


def test_row_sparse_table_update_equals_the_dense_path(ffk):
    """DLRM tables: scattering -lr * dOut straight into the table (ffk_embedding_sgd_update) must leave
    the same tables, MLP weights and outputs as gradient table + zero fill + dense SGD."""
    import importlib
    T = importlib.import_module("flexflow_train_b200.training")
    cfg = dict(batch=512, embedding_size=(2000,) * 4, sparse_feature_size=32, bag=3, mlp_bot=(8, 64, 32),
               mlp_top=(64, 32, 1))
    g = torch.Generator(device="cuda").manual_seed(4)
    dense = torch.rand(512, 8, device="cuda", generator=g)
    target = (torch.rand(512, 1, device="cuda", generator=g) > 0.5).float()
    nets = []
    for sparse in (False, True):
        net = T.DLRM(sparse_update=sparse, **cfg)
        opt = T.SGD(lr=0.3, weight_decay=0.0)
        net.dense.data.copy_(dense)
        for _ in range(4):
            T.dlrm_step(net, target, opt)
        torch.cuda.synchronize()
        nets.append(net)
    a, b = nets
    assert a.emb.dtables[0] is not None and b.emb.dtables[0] is None
    for ta, tb in zip(a.emb.tables, b.emb.tables):
        assert_close(tb, ta.cpu().numpy(), rtol=1e-5, what="table after row-sparse updates")
    assert_close(b.model.output.data, a.model.output.data.cpu().numpy(), rtol=1e-4, what="output")
