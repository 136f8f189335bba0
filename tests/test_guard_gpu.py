"""Out-of-bounds write check without a sanitizer: every tensor an entry point may write is carved
out of one slab pre-filled with a byte pattern, with 4 KiB guard bands before and after it and a
base that is 16-byte aligned but not 32-byte aligned (the weakest alignment the ABI promises
to accept). After the call the guard bands must still hold the pattern. Shapes are ragged on
purpose: edges are where vector tails, TMA boxes and tile epilogues go wrong."""
import pytest
import torch

pytestmark = pytest.mark.gpu

PATTERN = 0xA5
BAND = 4096


class Slab:
    def __init__(self, nbytes=256 << 20):
        self.buf = torch.full((nbytes,), PATTERN, dtype=torch.uint8, device="cuda")
        self.off = 0
        self.bands = []

    def take(self, shape, dtype, fill="rand", scale=1.0):
        numel = 1
        for s in shape:
            numel *= s
        nbytes = numel * torch.empty(0, dtype=dtype).element_size()
        base = self.buf.data_ptr()
        start = self.off + BAND
        start += (-(base + start)) % 32 + 16          # 16-byte aligned, not 32
        end = start + nbytes
        self.bands.append((self.off, start, end, end + BAND))
        self.off = end + BAND
        assert self.off + 64 <= self.buf.numel(), "slab too small for this test"
        t = self.buf[start:end].view(dtype).view(shape)
        if fill == "rand":
            if dtype in (torch.int32, torch.int64):
                t.zero_()
            else:
                t.copy_(((torch.rand(shape, device="cuda") * 2 - 1) * scale).to(dtype))
        elif fill == "zero":
            t.zero_()
        return t

    def check(self, what):
        torch.cuda.synchronize()
        for lo0, lo1, hi0, hi1 in self.bands:
            for a, b in ((lo0, lo1), (hi0, hi1)):
                band = self.buf[a:b]
                assert bool((band == PATTERN).all()), f"{what}: guard band [{a},{b}) was written"


@pytest.fixture()
def slab():
    return Slab()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(200, 136, 104), (129, 264, 520), (1, 8, 8), (300, 72, 1000), (257, 1032, 264)])
def test_linear_forward_backward_stay_in_bounds(K, slab, dtype, shape):
    batch, in_dim, out_dim = shape
    x = slab.take((batch, in_dim), dtype)
    w = slab.take((out_dim, in_dim), dtype, scale=0.05)
    b = slab.take((out_dim,), dtype)
    y = slab.take((batch, out_dim), dtype, "zero")
    dy = slab.take((batch, out_dim), dtype)
    dx = slab.take((batch, in_dim), dtype, "zero")
    dw = slab.take((out_dim, in_dim), dtype, "zero")
    db = slab.take((out_dim,), dtype, "zero")
    m = K.Linear.init_kernel(K.ACT_RELU, 1e-4, True, K.GEMM_AUTO, out_dim=out_dim)
    K.Linear.forward_kernel(None, m, x, y, w, b, in_dim, out_dim, batch)
    K.Linear.backward_kernel(None, m, x, dx, y, dy, w, dw, db, in_dim, out_dim, batch)
    slab.check(f"linear {dtype} {shape}")
    for t in (y, dx, dw, db):
        assert torch.isfinite(t.float()).all()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_batch_matmul_stays_in_bounds(K, slab, dtype):
    for (m_, n_, k_, batch) in [(72, 136, 40, 3), (264, 129, 72, 2), (8, 8, 8, 5)]:
        a = slab.take((batch, n_, k_), dtype)
        b = slab.take((batch, k_, m_), dtype)
        o = slab.take((batch, n_, m_), dtype, "zero")
        do = slab.take((batch, n_, m_), dtype)
        da = slab.take((batch, n_, k_), dtype, "zero")
        db = slab.take((batch, k_, m_), dtype, "zero")
        K.BatchMatmul.forward_kernel(None, o, a, b, m_, n_, k_, batch)
        K.BatchMatmul.backward_kernel(None, o, do, a, da, b, db, m_, n_, k_, batch)
        slab.check(f"bmm {dtype} {(m_, n_, k_, batch)}")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(7, 5), (33, 1000), (130, 4100), (3, 20004), (1000, 36)])
def test_row_kernels_stay_in_bounds(K, slab, dtype, rows, cols):
    x = slab.take((rows, cols), dtype)
    y = slab.take((rows, cols), dtype, "zero")
    K.Softmax.forward_kernel(None, x, y, rows, cols, 1)
    dx = slab.take((rows, cols), dtype, "zero")
    K.Softmax.backward_kernel(None, dx, x, y, rows, cols, 1, mode=1)
    if cols % 4 == 0:   # strided softmax over a middle dimension
        K.Softmax.forward_kernel(None, x, y, rows, cols // 4, 4)
    ln = K.LayerNorm.init_kernel(True, rows, cols, 1e-5)
    for name in ("mean", "rstd"):
        setattr(ln, name, slab.take((rows,), torch.float32, "zero"))
    gamma, beta = slab.take((cols,), dtype), slab.take((cols,), dtype)
    dgamma, dbeta = slab.take((cols,), dtype, "zero"), slab.take((cols,), dtype, "zero")
    K.LayerNorm.forward_kernel(None, ln, x, y, gamma, beta)
    K.LayerNorm.backward_kernel(None, ln, y, x, dx, gamma, dgamma, dbeta)
    ln2 = K.LayerNorm.init_kernel(True, rows, cols, 1e-5, workspace=False)
    ln2.mean, ln2.rstd = ln.mean, ln.rstd
    K.LayerNorm.backward_kernel(None, ln2, y, x, dx, gamma, dgamma, dbeta)
    slab.check(f"softmax/layernorm {dtype} {rows}x{cols}")
    assert torch.isfinite(y.float()).all() and torch.isfinite(dx.float()).all()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n", [1, 7, 1023, 65537, 1000003])
def test_elementwise_stays_in_bounds(K, slab, dtype, n):
    OT = K.OperatorType
    a, b = slab.take((n,), dtype), slab.take((n,), dtype)
    o = slab.take((n,), dtype, "zero")
    ga, gb = slab.take((n,), dtype, "zero"), slab.take((n,), dtype, "zero")
    for op in (OT.EW_ADD, OT.EW_MUL, OT.EW_MAX):
        K.ElementBinary.forward_kernel(None, op, a, b, o)
        K.ElementBinary.backward_kernel(None, op, o, a, b, ga, gb)
    for op, scalar in ((OT.RELU, 0.0), (OT.GELU, 0.0), (OT.SCALAR_MULTIPLY, 1.5), (OT.TANH, 0.0)):
        K.ElementUnary.forward_kernel(None, op, scalar, a, o)
        K.ElementUnary.backward_kernel(None, op, scalar, a, ga, o, b)
    d = K.Dropout.init_kernel(0.3, 5, n)
    K.Dropout.forward_kernel(None, d, a, o)
    K.Dropout.backward_kernel(None, d, o, ga)
    f32 = slab.take((n,), torch.float32, "zero")
    K.Cast.forward_kernel(None, a, f32)
    K.copy(None, o, a)
    K.accumulate(None, o, b)
    slab.check(f"elementwise {dtype} n={n}")


@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
@pytest.mark.parametrize("dim", [1, 3, 100, 260])
def test_embedding_gather_topk_stay_in_bounds(K, slab, idx_dtype, dim):
    entries, batch, bag = 517, 77, 3
    g = torch.Generator(device="cuda").manual_seed(5)
    table = slab.take((entries, dim), torch.float32)
    dtable = slab.take((entries, dim), torch.float32, "zero")
    for aggr, in_dim in ((K.AGGR_NONE, 1), (K.AGGR_SUM, bag), (K.AGGR_AVG, bag)):
        idx = slab.take((batch, in_dim), idx_dtype)
        idx.copy_(torch.randint(0, entries, (batch, in_dim), device="cuda", generator=g).to(idx_dtype))
        out = slab.take((batch, dim), torch.float32, "zero")
        K.Embedding.forward_kernel(None, idx, out, table, aggr, in_dim, dim, batch)
        K.Embedding.backward_kernel(None, idx, out, dtable, aggr, in_dim, dim, batch)
    # gather along dim 1 of [5, entries, dim] with [5, 9, dim] indices
    src = slab.take((5, 31, dim), torch.float32)
    gi = slab.take((5, 9, dim), idx_dtype)
    gi.copy_(torch.randint(0, 31, (5, 9, dim), device="cuda", generator=g).to(idx_dtype))
    go = slab.take((5, 9, dim), torch.float32, "zero")
    gsrc = slab.take((5, 31, dim), torch.float32, "zero")
    K.Gather.forward_kernel(None, 1, src, gi, go)
    K.Gather.backward_kernel(None, 1, go, gi, gsrc)
    # top-k over rows of length `dim`
    k = min(dim, 5)
    tv = slab.take((batch, k), torch.float32, "zero")
    ti = slab.take((batch, k), torch.int32)
    tin = slab.take((batch, dim), torch.float32)
    tg = slab.take((batch, dim), torch.float32, "zero")
    K.TopK.forward_kernel(None, tin, tv, ti, batch, dim, k)
    K.TopK.backward_kernel(None, tv, ti, tg, batch, dim, k)
    slab.check(f"embedding/gather/topk {idx_dtype} dim={dim}")


def test_shape_ops_stay_in_bounds(K, slab):
    x = slab.take((5, 7, 33, 9), torch.float32)
    for perm in ((0, 1, 3, 2), (3, 2, 1, 0), (0, 2, 1, 3), (1, 0, 2, 3)):
        shape = tuple(x.shape[p] for p in perm)
        y = slab.take(shape, torch.float32, "zero")
        m = K.Transpose.init_kernel(4, perm)
        K.Transpose.forward_kernel(None, m, x, y)
        gx = slab.take(tuple(x.shape), torch.float32, "zero")
        K.Transpose.backward_kernel(None, m, gx, y)
        assert torch.equal(y, x.permute(perm).contiguous())
    r = slab.take(tuple(x.shape), torch.float32, "zero")
    K.Reverse.forward_kernel(None, x, r, 5 * 7, 33, 9)
    assert torch.equal(r, x.flip(2))
    red = slab.take((5, 7, 1, 9), torch.float32, "zero")
    K.Reduce.forward_kernel(None, K.OperatorType.REDUCE_SUM, x, red)
    gx = slab.take(tuple(x.shape), torch.float32, "zero")
    K.Reduce.backward_kernel(None, K.OperatorType.REDUCE_SUM, red, gx)
    # concat / split along axis 1
    parts = [slab.take((4, c, 6), torch.float32) for c in (3, 1, 5)]
    cat = slab.take((4, 9, 6), torch.float32, "zero")
    K.Concat.forward_kernel(None, cat, parts, 1)
    assert torch.equal(cat, torch.cat(parts, 1))
    grads = [slab.take(tuple(p.shape), torch.float32, "zero") for p in parts]
    K.Concat.backward_kernel(None, cat, grads, 1)
    # pooling and convolution with ragged spatial sizes
    pin = slab.take((3, 5, 13, 11), torch.float32)
    oh, ow = (13 + 2 - 3) // 2 + 1, (11 + 2 - 3) // 2 + 1
    pout = slab.take((3, 5, oh, ow), torch.float32, "zero")
    pg = slab.take((3, 5, 13, 11), torch.float32, "zero")
    for kind in (K.POOL_MAX, K.POOL_AVG):
        pm = K.Pool2D.init_kernel(11, 13, 5, 3, ow, oh, 5, 3, 1, 1, 3, 3, 2, 2, kind)
        K.Pool2D.forward_kernel(None, pm, pin, pout)
        K.Pool2D.backward_kernel(None, pm, pin, pg, pout, pout)
    f = slab.take((6, 5, 3, 3), torch.float32, scale=0.1)
    cb = slab.take((6,), torch.float32)
    cout = slab.take((3, 6, oh, ow), torch.float32, "zero")
    cm = K.Conv2D.init_kernel(K.ACT_RELU, 3, 3, 1, 1, 1, 2, 2, (3, 5, 13, 11), (3, 6, oh, ow))
    K.Conv2D.forward_kernel(None, cm, pin, cout, f, cb)
    df, dcb = slab.take((6, 5, 3, 3), torch.float32, "zero"), slab.take((6,), torch.float32, "zero")
    dcout = slab.take((3, 6, oh, ow), torch.float32)
    K.Conv2D.backward_kernel(None, cm, pin, pg, cout, dcout, f, df, dcb)
    slab.check("shape ops")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_attention_stays_in_bounds(K, slab, dtype):
    n, heads, e, proj, sq, sk = 3, 2, 40, 24, 9, 13
    m = K.MultiHeadAttention.init_kernel(n, heads, e, e, e, proj, proj, proj, e, sq, sk, dtype=dtype)
    esz = torch.empty(0, dtype=dtype).element_size()
    m.reserveSpace = slab.take((max(m.reserveSpaceSize, 16),), torch.uint8, "zero")
    m.workSpace = slab.take((m.workSpace.numel(),), torch.uint8, "zero")
    q, k, v = slab.take((n, sq, e), dtype), slab.take((n, sk, e), dtype), slab.take((n, sk, e), dtype)
    w = slab.take((m.weightSize // esz,), dtype, scale=0.2)
    out = slab.take((n, sq, e), dtype, "zero")
    K.MultiHeadAttention.forward_kernel(None, m, q, k, v, w, out)
    dq, dk, dv = (slab.take(tuple(t.shape), dtype, "zero") for t in (q, k, v))
    dw = slab.take((m.weightSize // esz,), dtype, "zero")
    K.MultiHeadAttention.backward_kernel(None, m, q, dq, k, dk, v, dv, w, dw, out)
    slab.check(f"attention {dtype}")
    for t in (out, dq, dk, dv, dw):
        assert torch.isfinite(t.float()).all()
