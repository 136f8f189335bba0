"""The oracle's encoder step (tests/oracle_encoder.py, referee of the GPU encoder parity test and
CPU leg of bench.py) against PyTorch autograd on the CPU: same layer graph, attention without the
1/sqrt(d) scaling (cuDNN's smScalar = 1, attention_kernels.cu:114), MSE-average loss, SGD."""
import numpy as np
import torch

from oracle_encoder import OracleEncoder


def _torch_step(W, x, target, B, S, D, H, hd, FF, lr, wd, eps):
    P = [{k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in w.items()} for w in W]
    cur = torch.tensor(x, dtype=torch.float64).reshape(B, S, D)
    for p in P:
        sizes = [D * H * hd, D * H * hd, D * H * hd, H * hd * D]
        wq, wk, wv, wo = torch.split(p["mha"], sizes)
        q = torch.einsum("bse,ehd->bhsd", cur, wq.view(D, H, hd))
        k = torch.einsum("bse,ehd->bhsd", cur, wk.view(D, H, hd))
        v = torch.einsum("bse,ehd->bhsd", cur, wv.view(D, H, hd))
        pr = torch.softmax(q @ k.transpose(-1, -2), dim=-1)
        att = torch.einsum("bhsd,hde->bse", pr @ v, wo.view(H, hd, D))
        n1 = torch.nn.functional.layer_norm(cur + att, (D,), p["g1"], p["b1"], eps)
        h = torch.relu(n1 @ p["w1"].T + p["c1"])
        f = h @ p["w2"].T + p["c2"]
        cur = torch.nn.functional.layer_norm(n1 + f, (D,), p["g2"], p["b2"], eps)
    out = cur
    loss = ((out - torch.tensor(target, dtype=torch.float64).reshape(B, S, D)) ** 2).mean()
    loss.backward()
    new = [{k: (v - lr * (v.grad + wd * v)).detach().numpy() for k, v in p.items()} for p in P]
    return out.detach().numpy(), new


def test_oracle_encoder_step_matches_autograd(oracle):
    B, S, D, H, hd, FF = 2, 8, 16, 2, 12, 40
    W = OracleEncoder.random_weights(2, D, H, hd, FF, seed=3)
    rng = np.random.default_rng(0)
    for w in W:   # non-trivial affine parameters and biases
        for k in ("g1", "b1", "c1", "c2", "g2", "b2"):
            w[k] += rng.uniform(-0.3, 0.3, w[k].shape).astype(np.float32)
    x = rng.uniform(-1, 1, (B, S, D)).astype(np.float32)
    t = rng.uniform(-1, 1, (B, S, D)).astype(np.float32)
    ref_out, ref_w = _torch_step(W, x, t, B, S, D, H, hd, FF, 0.05, 1e-3, 1e-5)
    enc = OracleEncoder(oracle, W, B, S, D, H, hd, FF)
    out = enc.step(x, t, 0.05, 1e-3)
    np.testing.assert_allclose(out, ref_out, rtol=2e-5, atol=2e-6)
    for got, want in zip(W, ref_w):
        for k in got:
            np.testing.assert_allclose(got[k].reshape(-1), want[k].reshape(-1), rtol=1e-4, atol=1e-6, err_msg=k)
