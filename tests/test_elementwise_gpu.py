"""GPU parity: ElementUnary / ElementBinary / Cast / copy-like kernels / Dropout vs the oracle."""
import numpy as np
import pytest
import torch

from helpers import SEED_ACT, SEED_W, assert_close, assert_equal_bits, to_dev, to_np, uniform

pytestmark = pytest.mark.gpu


UNARY = ["SCALAR_MULTIPLY", "SCALAR_ADD", "SCALAR_SUB", "SCALAR_TRUE_DIV", "POW", "RELU", "SIGMOID",
         "TANH", "ELU", "EXP", "IDENTITY", "GELU", "RSQRT", "SIN", "COS"]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("name", UNARY)
def test_element_unary(K, oracle, name, dtype):
    op = getattr(K.OperatorType, name)
    n = 4096 * 3 + 5  # ragged tail after the 16-byte packets
    positive = name in ("POW", "RSQRT")
    xin = uniform((n,), SEED_ACT, 0.1 if positive else -1.0, 2.0 if positive else 1.0)
    scalar = {"POW": 1.5, "ELU": 1.0}.get(name, 0.75)
    x, xn = to_dev(xin, dtype)
    y = torch.empty_like(x)
    K.ElementUnary.forward_kernel(None, op, scalar, x, y)
    yref = np.zeros(n, np.float32)
    oracle.element_unary_forward(op, scalar, xn, yref, n)
    assert_close(y, yref, dtype, what=f"unary fwd {name}")
    dy, dyn = to_dev(uniform((n,), SEED_ACT + 1), dtype)
    dx, dxn = to_dev(uniform((n,), SEED_ACT + 2), dtype)
    yq, yn = to_dev(yref, dtype)
    K.ElementUnary.backward_kernel(None, op, scalar, x, dx, yq, dy)
    oracle.element_unary_backward(op, scalar, xn, dxn, yn, dyn, n)
    assert_close(dx, dxn, dtype, what=f"unary bwd {name}")


BINARY = ["EW_ADD", "EW_SUB", "EW_MUL", "EW_DIV", "EW_MAX", "EW_MIN"]
SHAPES = [
    ((4, 3, 8, 16), (4, 3, 8, 16)),     # same shape
    ((4, 3, 8, 16), (1, 3, 1, 1)),      # rhs broadcast (per-channel, cuDNN style)
    ((1, 1, 1, 16), (4, 3, 8, 16)),     # lhs broadcast (bias-like)
    ((5, 1, 7), (5, 6, 1)),             # both sides broadcast
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("name", BINARY)
@pytest.mark.parametrize("lshape,rshape", SHAPES)
def test_element_binary(K, oracle, name, lshape, rshape, dtype):
    op = getattr(K.OperatorType, name)
    oshape = tuple(max(a, b) for a, b in zip(lshape, rshape))
    l, ln = to_dev(uniform(lshape, SEED_ACT), dtype)
    rin = uniform(rshape, SEED_W)
    if name == "EW_DIV":
        rin = np.sign(rin) * (np.abs(rin) + 0.5)
    r, rn = to_dev(rin, dtype)
    o = torch.empty(oshape, dtype=dtype, device="cuda")
    K.ElementBinary.forward_kernel(None, op, l, r, o)
    oref = np.zeros(oshape, np.float32)
    oracle.element_binary_forward(op, ln, rn, oref, len(oshape), lshape, rshape, oshape)
    assert_close(o, oref, dtype, what=f"binary fwd {name} {lshape}{rshape}")
    dy, dyn = to_dev(uniform(oshape, SEED_ACT + 1), dtype)
    dl, dln = to_dev(uniform(lshape, SEED_ACT + 2), dtype)
    dr, drn = to_dev(uniform(rshape, SEED_ACT + 3), dtype)
    K.ElementBinary.backward_kernel(None, op, dy, l, r, dl, dr)
    oracle.element_binary_backward(op, dyn, ln, rn, dln, drn, len(oshape), lshape, rshape, oshape)
    assert_close(dl, dln, dtype, what=f"binary bwd lhs {name} {lshape}{rshape}")
    assert_close(dr, drn, dtype, what=f"binary bwd rhs {name} {lshape}{rshape}")


def test_element_binary_full_size_add_is_exact(K):
    """config[1] size 8192x4096 fp32: a+b must be the IEEE sum, element for element."""
    g = torch.Generator(device="cuda").manual_seed(SEED_ACT)
    a = torch.rand(8192, 4096, device="cuda", generator=g)
    b = torch.rand(8192, 4096, device="cuda", generator=g)
    o = torch.empty_like(a)
    K.ElementBinary.forward_kernel(None, K.OperatorType.EW_ADD, a, b, o)
    assert torch.equal(o, a + b)
    da, db = torch.ones_like(a), torch.ones_like(a)
    K.ElementBinary.backward_kernel(None, K.OperatorType.EW_ADD, o, a, b, da, db)
    assert torch.equal(da, 1 + o) and torch.equal(db, 1 + o)


@pytest.mark.parametrize("src,dst", [(torch.float32, torch.int32), (torch.int32, torch.float32),
                                     (torch.float32, torch.bfloat16), (torch.bfloat16, torch.float32),
                                     (torch.float32, torch.float16), (torch.int64, torch.float64),
                                     (torch.float64, torch.float32), (torch.float32, torch.int64)])
def test_cast(K, src, dst):
    n = 1000
    base = (uniform((n,), SEED_ACT) * 100)
    x = torch.from_numpy(base).cuda().to(src)
    y = torch.zeros(n, dtype=dst, device="cuda")
    K.Cast.forward_kernel(None, x, y)
    assert torch.equal(y, x.to(dst)), "cast must match a C static_cast / round-to-nearest"
    acc0 = torch.arange(n, device="cuda").to(dst)
    acc = acc0.clone()
    K.Cast.backward_kernel(None, x, acc)
    if dst in (torch.bfloat16, torch.float16):
        expect = (x.float() + acc0.float()).to(dst)
    else:
        expect = x.to(dst) + acc0
    assert torch.equal(acc, expect)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int32, torch.int64, torch.bool, torch.float64])
@pytest.mark.parametrize("n", [0, 1, 100, 10000, 100 * 100, 4097])
def test_copy_family_is_bit_exact(K, dtype, n):
    """Flat / Reshape / Combine / Repartition / Replicate fwd, Reduction bwd: byte copies."""
    g = torch.Generator(device="cuda").manual_seed(n + 1)
    raw = torch.randint(0, 255, (max(n, 1) * 8,), dtype=torch.uint8, device="cuda", generator=g)
    if dtype == torch.bool:
        src = (raw[:n] & 1).bool()
    else:
        src = raw[: n * torch.empty(0, dtype=dtype).element_size()].view(dtype)
    for op in (K.Flat, K.Reshape, K.Combine, K.Repartition, K.Replicate):
        dst = torch.zeros_like(src)
        op.forward_kernel(None, src, dst)
        assert_equal_bits(dst, src, what=op.__name__)
    dst = torch.zeros_like(src)
    K.Reduction.backward_kernel(None, dst, src)
    assert_equal_bits(dst, src, what="Reduction.backward")


def test_copy_unaligned_pointers(K):
    base = torch.arange(4096, dtype=torch.uint8, device="cuda")
    src = base[3:3 + 1001]
    dst_base = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    dst = dst_base[5:5 + 1001]
    import flexflow_train_b200.kernels as kk
    kk._call("ffk_copy", None, dst.data_ptr(), src.data_ptr(), 1001, kk.BOOL)
    assert torch.equal(dst, src) and dst_base[:5].sum() == 0 and dst_base[5 + 1001:].sum() == 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_accumulate_family(K, oracle, dtype):
    n = 12345
    for op, order in ((K.Flat, "flat"), (K.Reshape, "std"), (K.Combine, "combine"), (K.Repartition, "std")):
        g, gn = to_dev(uniform((n,), SEED_ACT), dtype)
        acc, accn = to_dev(uniform((n,), SEED_W), dtype)
        if order == "flat":
            op.backward_kernel(None, None, acc, g)
        elif order == "combine":
            op.backward_kernel(None, g, acc)
        else:
            op.backward_kernel(None, acc, g)
        oracle.accumulate(accn, gn, n)
        assert_close(acc, accn, dtype, what=op.__name__)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n,reps", [(100, 10), (100000, 5), (4099, 3), (1, 1)])
def test_replica_sums(K, oracle, dtype, n, reps):
    src, srcn = to_dev(uniform((reps, n), SEED_ACT), dtype)
    out = torch.empty(n, dtype=dtype, device="cuda")
    K.Reduction.forward_kernel(None, src, out, reps)
    ref = np.zeros(n, np.float32)
    oracle.sum_replicas(ref, srcn, n, reps, 0)
    assert_close(out, ref, dtype, what="Reduction.forward")
    acc, accn = to_dev(uniform((n,), SEED_W), dtype)
    K.Replicate.backward_kernel(None, acc, src, reps)
    oracle.sum_replicas(accn, srcn, n, reps, 1)
    assert_close(acc, accn, dtype, what="Replicate.backward")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [10 * 10, 32 * 1000, 32 * 1000 + 17, 5])
def test_dropout(K, oracle, dtype, n):
    rate, seed = 0.1, 12345  # reference fixture: lib/kernels/test/src/test_dropout_kernel.cc:9-10
    x, xn = to_dev(uniform((n,), SEED_ACT, 0.5, 1.5), dtype)
    y = torch.empty_like(x)
    m = K.Dropout.init_kernel(rate, seed, n)
    K.Dropout.forward_kernel(None, m, x, y)
    mask = m.mask.cpu().numpy().view(np.uint32)
    yref = np.zeros(n, np.float32)
    oracle.dropout_apply(xn, yref, mask, n, rate)
    assert_close(y, yref, dtype, what="dropout fwd given mask")
    kept = np.unpackbits(mask.view(np.uint8), bitorder="little")[:n]
    assert ((to_np(y) != 0) == (kept == 1)).all(), "mask bits must describe the output"
    if n >= 1000:
        assert abs(kept.mean() - (1 - rate)) < 0.02, "keep rate must be 1 - rate"
    dy, dyn = to_dev(uniform((n,), SEED_ACT + 1), dtype)
    dx = torch.full_like(dy, 9.0)  # overwritten, not accumulated (dropout_kernels.cu:86-93)
    K.Dropout.backward_kernel(None, m, dy, dx)
    dref = np.zeros(n, np.float32)
    oracle.dropout_apply(dyn, dref, mask, n, rate)
    assert_close(dx, dref, dtype, what="dropout bwd")
    # same (seed, offset) => same mask regardless of alignment path; new offset => new mask
    m2 = K.Dropout.init_kernel(rate, seed, n)
    y2 = torch.empty_like(x)
    K.Dropout.forward_kernel(None, m2, x, y2)
    assert torch.equal(m2.mask, m.mask)
    K.Dropout.forward_kernel(None, m2, x, y2)
    if n >= 1000:
        assert not torch.equal(m2.mask, m.mask)


def test_dropout_mask_independent_of_vector_path(K):
    n = 32 * 64 + 9
    base = torch.ones(n + 1, device="cuda")
    m1 = K.Dropout.init_kernel(0.3, 7, n)
    m2 = K.Dropout.init_kernel(0.3, 7, n)
    K.Dropout.forward_kernel(None, m1, base[:n].clone(), torch.empty(n, device="cuda"))
    import flexflow_train_b200.kernels as kk
    out = torch.empty(n + 1, device="cuda")
    kk._call("ffk_dropout_forward", None, base[1:].data_ptr(), out[1:].data_ptr(), m2.mask.data_ptr(),
             n, 0.3, 7, 0, kk.FLOAT)  # 4-byte aligned only -> generic path
    assert torch.equal(m1.mask, m2.mask)
