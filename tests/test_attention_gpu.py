"""GPU parity: MultiHeadAttention through the C ABI against (a) the golden vectors produced by
cuDNN's own multi-head attention as the reference configures it (tests/golden/mha_*.npz,
TF32 arithmetic inside cuDNN -> 2e-3 of scale) and (b) the fp64-accumulating CPU oracle
(rtol 1e-4 fp32, 2e-2 bf16)."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from helpers import assert_close, assert_close_l2, to_dev, uniform

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
KEYS = ("N", "H", "qSize", "kSize", "vSize", "qProj", "kProj", "vProj", "oProj", "qoSeq", "kvSeq")


def run_mha(K, dims, q, k, v, w, dout, dtype, path=0, skip_kv_grads=False):
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    m = K.MultiHeadAttention.init_kernel(*dims, dtype=dtype, path=path)
    assert m.weightSize == w.numel() * w.element_size()
    out = torch.zeros(N, Sq, oP, dtype=dtype, device="cuda")
    K.MultiHeadAttention.forward_kernel(None, m, q, k, v, w, out)
    dq = torch.full_like(q, 7.0)  # must be overwritten
    dk = None if skip_kv_grads else torch.full_like(k, 7.0)
    dv = None if skip_kv_grads else torch.full_like(v, 7.0)
    dw = torch.ones_like(w)       # must be accumulated into
    K.MultiHeadAttention.backward_kernel(None, m, q, dq, k, dk, v, dv, w, dw, dout)
    return out, dq, dk, dv, dw


@pytest.mark.parametrize("case", ["fixture", "ragged", "cross"])
def test_mha_fp32_against_cudnn_golden_vectors(K, case):
    z = np.load(GOLD / f"mha_{case}.npz")
    c = json.loads(str(z["layout"]))["config"]
    dims = [c[k] for k in KEYS]
    t = lambda n: torch.from_numpy(np.ascontiguousarray(z[n])).cuda()
    out, dq, dk, dv, dw = run_mha(K, dims, t("q"), t("k"), t("v"), t("w"), t("dout"), torch.float32)
    for got, name, off in ((out, "out", 0), (dq, "dq", 0), (dk, "dk", 0), (dv, "dv", 0), (dw, "dw", 1)):
        ref = z[name] + off
        err = np.abs(got.cpu().numpy().reshape(-1) - ref).max()
        assert err <= 2e-3 * np.abs(z[name]).max(), f"{case} {name}: {err}"


@pytest.mark.parametrize("dtype,path", [(torch.float32, 0), (torch.bfloat16, 0), (torch.bfloat16, 1)])
@pytest.mark.parametrize("dims", [[10, 4, 64, 64, 64, 64, 64, 64, 64, 20, 20],   # the reference's fixture
                                  [2, 2, 32, 48, 40, 16, 16, 24, 56, 33, 65],
                                  [1, 8, 128, 128, 128, 32, 32, 32, 128, 128, 128]])
def test_mha_against_oracle(K, oracle, dims, dtype, path):
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    if path == 1 and any(x % 8 for x in (Sk, qP, vP, qS, kS, vS, oP)):
        pytest.skip("forcing the tensor-core path needs 16-byte aligned pitches (AUTO falls back)")
    q, qn = to_dev(uniform((N, Sq, qS), 1234), dtype)
    k, kn = to_dev(uniform((N, Sk, kS), 1235), dtype)
    v, vn = to_dev(uniform((N, Sk, vS), 1236), dtype)
    nw = qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP
    w, wn = to_dev(uniform((nw,), 4321) * (2.0 / np.sqrt(qS)), dtype)
    dout, doutn = to_dev(uniform((N, Sq, oP), 1237), dtype)
    out, dq, dk, dv, dw = run_mha(K, dims, q, k, v, w, dout, dtype, path)
    ref = np.zeros((N, Sq, oP), np.float32)
    oracle.mha_forward(dims, qn, kn, vn, wn, ref)
    rdq, rdk, rdv, rdw = np.zeros_like(qn), np.zeros_like(kn), np.zeros_like(vn), np.ones_like(wn)
    oracle.mha_backward(dims, qn, rdq, kn, rdk, vn, rdv, wn, rdw, doutn)
    pairs = ((out, ref, "out"), (dq, rdq, "dq"), (dk, rdk, "dk"), (dv, rdv, "dv"), (dw, rdw, "dw"))
    for got, want, name in pairs:
        if dtype == torch.float32:
            assert_close(got, want, dtype, what="mha " + name)
        else:
            # projections, probabilities and their gradients are stored in bf16 between the
            # seven GEMMs of the chain: see helpers.assert_close_l2
            assert_close_l2(got, want, what="mha " + name)


def test_mha_null_key_value_grads_and_bad_config(K, ffk):
    dims = [2, 2, 16, 16, 16, 8, 8, 8, 16, 4, 4]
    q, _ = to_dev(uniform((2, 4, 16), 1))
    w, _ = to_dev(uniform((16 * 16 * 3 + 16 * 16,), 2))
    dout, _ = to_dev(uniform((2, 4, 16), 3))
    out, dq, dk, dv, dw = run_mha(K, dims, q, q, q, w, dout, torch.float32, skip_kv_grads=True)
    assert dk is None and dv is None and torch.isfinite(dq).all()
    with pytest.raises(ffk.FFKError):
        K.MultiHeadAttention.init_kernel(2, 2, 16, 16, 16, 8, 4, 8, 16, 4, 4)  # qProj != kProj


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("dims", [[2, 2, 64, 64, 64, 32, 32, 32, 64, 512, 512],
                                  [1, 3, 64, 64, 64, 64, 64, 64, 64, 300, 512],   # ragged query block
                                  [3, 2, 32, 32, 32, 256, 256, 256, 32, 256, 512]])  # the encoder's head dim
def test_mha_512_keys_scores_normalised_in_tmem(K, oracle, dims, dtype):
    """512-key rows of 16-bit tensors take the fused score+softmax kernel (probabilities are
    normalised in TMEM and stored once). Checked against the oracle and against the unfused
    CUDA-core path of the same entry point."""
    N, H, qS, kS, vS, qP, kP, vP, oP, Sq, Sk = dims
    q, qn = to_dev(uniform((N, Sq, qS), 1234), dtype)
    k, kn = to_dev(uniform((N, Sk, kS), 1235), dtype)
    v, vn = to_dev(uniform((N, Sk, vS), 1236), dtype)
    nw = qS * H * qP + kS * H * kP + vS * H * vP + H * vP * oP
    w, wn = to_dev(uniform((nw,), 4321) * (2.0 / np.sqrt(qS)), dtype)
    dout, doutn = to_dev(uniform((N, Sq, oP), 1237), dtype)
    out, dq, dk, dv, dw = run_mha(K, dims, q, k, v, w, dout, dtype, 0)
    ref = np.zeros((N, Sq, oP), np.float32)
    oracle.mha_forward(dims, qn, kn, vn, wn, ref)
    rdq, rdk, rdv, rdw = np.zeros_like(qn), np.zeros_like(kn), np.zeros_like(vn), np.ones_like(wn)
    oracle.mha_backward(dims, qn, rdq, kn, rdk, vn, rdv, wn, rdw, doutn)
    for got, want, name in ((out, ref, "out"), (dq, rdq, "dq"), (dk, rdk, "dk"), (dv, rdv, "dv"), (dw, rdw, "dw")):
        assert_close_l2(got, want, what="mha512 " + name)
    out2, _, _, _, _ = run_mha(K, dims, q, k, v, w, dout, dtype, K.GEMM_SIMT)
    assert_close_l2(out, out2.float().cpu().numpy(), rel=2e-2, what="fused vs unfused out")
