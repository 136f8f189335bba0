"""Shared helpers for the parity tests: seeded synthetic inputs (SURVEY §8d seeds), moving
them to the GPU in a given dtype, and the tolerance rule.

Tolerances (BASELINE.json north_star): bit-exact for copy / index paths; rtol 1e-4 for fp32
and 2e-2 for bf16/half elsewhere. Every element must satisfy
    |got - ref| <= atol + rtol * |ref|,   atol = rtol * mean(|ref|)
(the absolute term covers results that cancel to ~0, where a pure relative bound is
meaningless)."""
import numpy as np
import torch

SEED_ACT, SEED_W, SEED_IDX = 1234, 4321, 99
RTOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2, torch.float16: 2e-2}


def uniform(shape, seed, lo=-1.0, hi=1.0):
    rng = np.random.default_rng(seed)
    return rng.uniform(lo, hi, size=shape).astype(np.float32)


def to_dev(a, dtype=torch.float32):
    """numpy fp32 -> CUDA tensor of `dtype`; also returns the fp32 numpy view of what the
    GPU actually holds (inputs rounded to the storage dtype), which is what the oracle gets."""
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda().to(dtype).contiguous()
    return t, t.float().cpu().numpy()


def to_np(t):
    return t.detach().float().cpu().numpy()


def assert_close(got, ref, dtype=torch.float32, rtol=None, what=""):
    got = to_np(got) if isinstance(got, torch.Tensor) else np.asarray(got, dtype=np.float32)
    ref = np.asarray(ref, dtype=np.float32).reshape(got.shape)
    rtol = RTOL[dtype] if rtol is None else rtol
    atol = rtol * float(np.mean(np.abs(ref))) + 1e-30
    err = np.abs(got - ref)
    bound = atol + rtol * np.abs(ref)
    bad = err > bound
    if bad.any():
        i = np.unravel_index(np.argmax(err - bound), err.shape)
        raise AssertionError(
            f"{what}: {bad.sum()} of {bad.size} elements outside rtol={rtol} atol={atol:.3e}; "
            f"worst at {i}: got {got[i]!r} ref {ref[i]!r}")


def assert_equal_bits(got, ref, what=""):
    g = got.detach().cpu().contiguous().view(torch.uint8).numpy() if isinstance(got, torch.Tensor) else np.asarray(got).view(np.uint8)
    r = ref.detach().cpu().contiguous().view(torch.uint8).numpy() if isinstance(ref, torch.Tensor) else np.ascontiguousarray(ref).view(np.uint8)
    assert g.shape == r.shape and (g == r).all(), f"{what}: not bit-identical"


def assert_close_l2(got, ref, rel=2e-2, max_rel=0.15, what=""):
    """Bound for chains of 16-bit GEMMs (attention): every intermediate is rounded to 16 bits,
    so single elements that cancel can miss an elementwise 2e-2; the tensor as a whole must
    still agree to 2e-2 in relative L2 norm and no element may be off by more than
    `max_rel` of the tensor's largest magnitude."""
    got = to_np(got) if isinstance(got, torch.Tensor) else np.asarray(got, dtype=np.float32)
    ref = np.asarray(ref, dtype=np.float32).reshape(got.shape)
    l2 = np.linalg.norm((got - ref).ravel()) / max(np.linalg.norm(ref.ravel()), 1e-30)
    worst = np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)
    assert l2 <= rel, f"{what}: relative L2 error {l2:.3e} > {rel}"
    assert worst <= max_rel, f"{what}: worst element off by {worst:.3e} of the tensor scale"
