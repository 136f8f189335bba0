"""GPU parity for the K-cut schedule of the CTA-pair GEMM (gemm_sm100.cu `sk_begin`): when a
tile count leaves the last wave of CTA pairs mostly idle and scratch is registered, the tiles of
that wave are cut along K, partial accumulators are parked in the scratch and added by the pair
that owns the tile. Checked against the CPU oracle (Linear forward / backward at a shape where all
three products are cut), against fp64 products for every operand layout with ragged edges, against
the same launch without scratch (whole tiles), and for run-to-run bit equality.

The reference has no counterpart (its products are cublasGemmEx calls, linear_kernels.cu:127-149);
the contract is the one of every other 16-bit product: rtol 2e-2 against the oracle."""
import os

import numpy as np
import pytest
import torch

from helpers import assert_close
from test_dense_gpu import _linear_case

pytestmark = pytest.mark.gpu


@pytest.fixture
def scratch(K):
    K.set_gemm_workspace(32 << 20)
    yield
    K.set_gemm_workspace(None)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_linear_with_k_cut_products_matches_the_oracle(K, oracle, scratch, dtype):
    # forward 4096x512x512 and dgrad: 32 tiles of 8 k-blocks -> 2 pieces of 4 k-blocks;
    # wgrad 512x512x4096: 4 tiles of 64 k-blocks -> 16 pieces (pairs) per tile
    _linear_case(K, oracle, dtype, 4096, 512, 512, "gelu", True)
    _linear_case(K, oracle, dtype, 4096, 512, 512, "none", False, l2=1e-3)


def _operands(M, N, Kd, a_mn, b_mn, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = ((torch.rand((Kd, M) if a_mn else (M, Kd), device="cuda", generator=g) * 2 - 1)).to(dtype)
    b = ((torch.rand((Kd, N) if b_mn else (N, Kd), device="cuda", generator=g) * 2 - 1)).to(dtype)
    A = a.t() if a_mn else a
    B = b.t() if b_mn else b
    return a, b, A, B


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(1000, 1304, 4104), (2000, 1304, 1032)], ids=["pieces", "heads-tails"])
def test_k_cut_gemm_every_layout_ragged(K, a_mn, b_mn, out_dtype, shape):
    """Ragged tiles on 74 pairs, both forms of the cut: 1000 x 1304 x 4104 = 24 tiles of 65
    k-blocks, 3 pieces of 22 each; 2000 x 1304 x 1032 = 48 tiles of 17 k-blocks, heads of 12 by the
    tiles' own pairs and tails of 5 taken two at a time by the other 26 pairs. alpha and the
    accumulate flag are applied once, by the owner."""
    M, N, Kd = shape
    a, b, A, B = _operands(M, N, Kd, a_mn, b_mn, torch.bfloat16, 7)
    g = torch.Generator(device="cuda").manual_seed(8)
    d0 = (torch.rand(M, N, device="cuda", generator=g) * 4 - 2).to(out_dtype)
    ref = (0.5 * (A.double() @ B.double().t()) + d0.double()).float().cpu().numpy()

    def run():
        d = d0.clone()
        K.Gemm.run(a, b, d, M, N, Kd, a_mn=a_mn, b_mn=b_mn, alpha=0.5, accumulate=True, path=K.GEMM_TENSOR)
        return d

    whole = run()                       # no scratch: whole tiles
    K.set_gemm_workspace(32 << 20)
    try:
        cut1, cut2 = run(), run()
    finally:
        K.set_gemm_workspace(None)
    assert torch.equal(cut1, cut2), "K-cut product is not reproducible"
    rtol = 2e-2 if out_dtype == torch.bfloat16 else 2e-3   # fp32 out: bf16 inputs are exact, fp32 accumulation order differs
    assert_close(cut1, ref, rtol=rtol, what=f"k-cut gemm {a_mn}{b_mn} {out_dtype}")
    assert_close(cut1, whole.float().cpu().numpy(), rtol=rtol, what="k-cut vs whole tiles")


def test_k_cut_switch_restores_whole_tiles(K, scratch):
    """ffk_set_gemm_kcut(0): with scratch registered the product must run whole tiles again, i.e.
    give the bits of the launch without scratch; switching it back on gives the cut result."""
    M, N, Kd = 768, 768, 8192
    a, b, _, _ = _operands(M, N, Kd, True, True, torch.bfloat16, 21)

    def run():
        d = torch.zeros(M, N, dtype=torch.float32, device="cuda")
        K.Gemm.run(a, b, d, M, N, Kd, a_mn=True, b_mn=True, path=K.GEMM_TENSOR)
        return d

    cut = run()
    K.set_gemm_kcut(False)
    try:
        off = run()
    finally:
        K.set_gemm_kcut(True)
    again = run()
    K.set_gemm_workspace(None)
    whole = run()
    assert torch.equal(off, whole), "switch off: not the whole-tile result"
    assert torch.equal(cut, again), "switch back on: not the cut result"
    assert_close(cut, whole.cpu().numpy(), rtol=1e-4, what="cut vs whole, fp32 output")


SHAPES = [  # M, N, K, a_mn, b_mn, what
    (4096, 4096, 8192, True, True, "configs[1] wgrad: 256 tiles, 3.46 waves"),
    (4096, 1024, 32768, True, True, "encoder wgrad: 64 tiles on 74 pairs, left whole (13.5 % idle: measured slower cut)"),
    (1024, 4096, 32768, True, True, "encoder wgrad (feed-forward 2), left whole"),
    (2048, 1280, 32768, True, True, "40 tiles: heads and tails"),
    (768, 768, 32768, True, True, "BERT-base wgrad: 9 tiles"),
    (3072, 768, 32768, True, True, "BERT-base wgrad (feed-forward): 36 tiles"),
    (8192, 4096, 4096, False, False, "configs[1] forward: 512 tiles, left alone"),
]


@pytest.mark.parametrize("shape", SHAPES[:6], ids=lambda s: f"{s[0]}x{s[1]}x{s[2]}")
def test_k_cut_gemm_training_shapes(K, scratch, shape):
    M, N, Kd, a_mn, b_mn, _ = shape
    a, b, A, B = _operands(M, N, Kd, a_mn, b_mn, torch.bfloat16, 11)
    d0 = torch.full((M, N), 0.25, dtype=torch.bfloat16, device="cuda")
    d = d0.clone()
    K.Gemm.run(a, b, d, M, N, Kd, a_mn=a_mn, b_mn=b_mn, accumulate=True, path=K.GEMM_TENSOR)
    ref = (A.float() @ B.float().t() + 0.25).cpu().numpy()
    assert_close(d, ref, torch.bfloat16, what=f"k-cut gemm {M}x{N}x{Kd}")


@pytest.mark.skipif(not os.environ.get("FFK_STREAMK_TIMING"), reason="timing probe (FFK_STREAMK_TIMING=1)")
def test_k_cut_timing(K):
    lines = []
    for (M, N, Kd, a_mn, b_mn, what) in SHAPES:
        a, b, _, _ = _operands(M, N, Kd, a_mn, b_mn, torch.bfloat16, 3)
        # rotate over enough operand sets to exceed 2 x L2
        per_set = (a.numel() + b.numel() + M * N) * 2
        nset = max(2, int(np.ceil(2 * 130e6 / per_set)) + 1)
        sets = [(a.clone(), b.clone(), torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")) for _ in range(nset)]
        res = {}
        for mode in ("whole", "cut"):
            K.set_gemm_workspace(32 << 20 if mode == "cut" else None)
            iters = 20
            for i in range(3):
                x, y, d = sets[i % nset]
                K.Gemm.run(x, y, d, M, N, Kd, a_mn=a_mn, b_mn=b_mn, accumulate=True, path=K.GEMM_TENSOR)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(iters):
                x, y, d = sets[i % nset]
                K.Gemm.run(x, y, d, M, N, Kd, a_mn=a_mn, b_mn=b_mn, accumulate=True, path=K.GEMM_TENSOR)
            e1.record()
            torch.cuda.synchronize()
            res[mode] = e0.elapsed_time(e1) / iters * 1e3
        K.set_gemm_workspace(None)
        tf = 2.0 * M * N * Kd / 1e6
        lines.append(f"{M}x{N}x{Kd} a_mn={int(a_mn)} b_mn={int(b_mn)} | whole {res['whole']:.1f} us {tf / res['whole']:.0f} TF/s | "
                     f"cut {res['cut']:.1f} us {tf / res['cut']:.0f} TF/s | {what}")
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/r02_streamk_timing.txt", "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines))
