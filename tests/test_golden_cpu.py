"""CPU: the oracle against the committed golden vectors.

tests/golden/mha_*.npz were produced by the reference's own arithmetic provider: cuDNN's
legacy multi-head attention, configured exactly as lib/kernels/src/cuda/ops/attention_kernels.cu
configures it, run on a B200 (oracle/cudnn_ref/mha_golden.cu, tools/make_mha_golden.py).
cuDNN evaluates those products with TF32 tensor-op math (10-bit mantissa), so the agreement
bound against them is 2e-3 of the tensor's scale; the oracle itself accumulates in fp64.

tests/golden/copy_like.json restates the exact-equality cases of the reference's kernel tests."""
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
CASES = ["fixture", "ragged", "cross"]


def load(case):
    z = np.load(GOLD / f"mha_{case}.npz")
    layout = json.loads(str(z["layout"]))
    c = layout["config"]
    dims = [c[k] for k in ("N", "H", "qSize", "kSize", "vSize", "qProj", "kProj", "vProj", "oProj", "qoSeq", "kvSeq")]
    return z, layout, dims


def close_tf32(got, ref, what):
    scale = np.abs(ref).max()
    err = np.abs(got - ref).max()
    assert err <= 2e-3 * scale, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("case", CASES)
def test_weight_layout_matches_cudnn_report(case):
    _, layout, dims = load(case)
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    kinds = layout["kinds"]
    assert layout["weightSizeBytes"] == 4 * (qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP)
    assert kinds["Q"] == {"offset_elems": 0, "dims": [H, qP, qS], "strides": [qP, 1, H * qP]}
    assert kinds["K"] == {"offset_elems": qS * H * qP, "dims": [H, kP, kS], "strides": [kP, 1, H * kP]}
    assert kinds["V"] == {"offset_elems": qS * H * qP + kS * H * kP, "dims": [H, vP, vS], "strides": [vP, 1, H * vP]}
    assert kinds["O"] == {"offset_elems": qS * H * qP + kS * H * kP + vS * H * vP, "dims": [H, oP, vP],
                          "strides": [vP * oP, 1, oP]}


@pytest.mark.parametrize("case", CASES)
def test_oracle_mha_against_cudnn_vectors(oracle, case):
    z, _, dims = load(case)
    q, k, v, w, dout = (np.ascontiguousarray(z[n]) for n in ("q", "k", "v", "w", "dout"))
    out = np.zeros_like(z["out"])
    oracle.mha_forward(dims, q, k, v, w, out)
    close_tf32(out, z["out"], "forward")
    dq, dk, dv, dw = (np.zeros_like(z[n]) for n in ("dq", "dk", "dv", "dw"))
    oracle.mha_backward(dims, q, dq, k, dk, v, dv, w, dw, dout)
    close_tf32(dq, z["dq"], "dq")
    close_tf32(dk, z["dk"], "dk")
    close_tf32(dv, z["dv"], "dv")
    close_tf32(dw, z["dw"], "dw")


def test_oracle_mha_against_torch(oracle):
    import torch

    dims = [2, 3, 10, 12, 14, 4, 4, 6, 8, 5, 7]
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    rng = np.random.default_rng(0)
    u = lambda *s: rng.uniform(-1, 1, s).astype(np.float32)
    q, k, v, dout = u(N, Sq, qS), u(N, Sk, kS), u(N, Sk, vS), u(N, Sq, oP)
    w = u(qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP) * 0.3
    tq, tk, tv, tw = (torch.from_numpy(a.copy()).double().requires_grad_() for a in (q, k, v, w))
    o = 0
    wq = tw[o:o + qS * H * qP].view(qS, H, qP); o += qS * H * qP
    wk = tw[o:o + kS * H * kP].view(kS, H, kP); o += kS * H * kP
    wv = tw[o:o + vS * H * vP].view(vS, H, vP); o += vS * H * vP
    wo = tw[o:].view(H, vP, oP)
    qp = torch.einsum("nsi,ihp->nhsp", tq, wq)
    kp = torch.einsum("nti,ihp->nhtp", tk, wk)
    vp = torch.einsum("nti,ihp->nhtp", tv, wv)
    p = torch.softmax(torch.einsum("nhsp,nhtp->nhst", qp, kp), -1)
    out_t = torch.einsum("nshp,hpo->nso", torch.einsum("nhst,nhtp->nshp", p, vp), wo)
    out_t.backward(torch.from_numpy(dout).double())
    out = np.zeros((N, Sq, oP), np.float32)
    oracle.mha_forward(dims, q, k, v, w, out)
    np.testing.assert_allclose(out, out_t.detach().numpy(), rtol=1e-5, atol=1e-6)
    dq, dk, dv, dw = np.full_like(q, 3.0), np.zeros_like(k), np.zeros_like(v), np.ones_like(w)
    oracle.mha_backward(dims, q, dq, k, dk, v, dv, w, dw, dout)
    np.testing.assert_allclose(dq, tq.grad.numpy(), rtol=1e-5, atol=1e-6)  # overwritten
    np.testing.assert_allclose(dk, tk.grad.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(dv, tv.grad.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(dw, 1 + tw.grad.numpy(), rtol=1e-5, atol=1e-5)  # accumulated
