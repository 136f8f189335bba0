"""Seeded cases run three ways — by the reference's OWN CUDA kernels (oracle/_ref/libref_kernels.so,
built from the reference checkout by oracle/ref_kernels; `run_ref`), by the CPU oracle
(`run_oracle`) and by libffk.so (`run_ffk`). tools/make_ref_golden.py runs the first on a B200 and
stores the outputs as tests/golden/ref_kernels.npz; the tests compare the other two against that
file (and, where libref_kernels.so travelled to the GPU box, against a live run). Test
infrastructure only. Inputs are U(-1,1) from numpy.default_rng(seed), SURVEY §8(d)'s convention.

A few outputs are where this repository deliberately departs from the reference's arithmetic
(SURVEY App. A). They are listed in DEVIATIONS with the reason: the tests require
the oracle / kernels to DIFFER from the golden there in the documented way, and to agree with
`intended` values computed here."""
import ctypes
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
REF_SO = ROOT / "oracle" / "_ref" / "libref_kernels.so"

# harness.cu's private OperatorType order -> kernels.py / reference OperatorType values
REF_UNARY = {"EXP": 0, "IDENTITY": 1, "GELU": 6, "RSQRT": 7, "SIN": 8, "COS": 9,
             "SCALAR_MULTIPLY": 10, "SCALAR_ADD": 11, "SCALAR_SUB": 12, "SCALAR_TRUE_DIV": 13, "POW": 14}
FF_UNARY = {"EXP": 56, "IDENTITY": 14, "GELU": 67, "RSQRT": 70, "SIN": 61, "COS": 62,
            "SCALAR_MULTIPLY": 8, "SCALAR_ADD": 9, "SCALAR_SUB": 12, "SCALAR_TRUE_DIV": 11, "POW": 71}
UNARY_OPS = list(REF_UNARY)

DEVIATIONS = {
    "unary_GELU_dx": "the reference's GELU derivative drops a factor (element_unary_kernels.cu:226-232); "
                     "the oracle and the kernels use the true derivative of x * Phi(x)",
    "emb_avg_i32_out": "the reference rescales inside the bag loop (embedding_kernels.cu:108-113): mean over the bag here",
    "emb_avg_i64_out": "same as emb_avg_i32_out",
}


def u(seed, *shape, lo=-1.0, hi=1.0):
    return np.random.default_rng(seed).uniform(lo, hi, shape).astype(np.float32)


def inputs():
    """Every case's inputs, by name."""
    I = {}
    for tag, (M, N) in (("ln10", (10, 10)), ("ln600", (600, 96))):      # M < 512 and M >= 512 column paths
        I[tag] = dict(M=M, N=N, eps=1e-5, x=u(1, M, N), gamma=u(2, N) + 1, beta=u(3, N), dy=u(4, M, N))
    rng = np.random.default_rng(99)
    I["emb"] = dict(entries=50, dim=16, batch=20, bag=3, table=u(5, 50, 16), dout=u(6, 20, 16),
                    idx=rng.integers(0, 50, (20, 3)), idx1=rng.integers(0, 50, 20))
    I["loss"] = dict(p=u(7, 8, 10, lo=0.0), y=u(8, 8, 10, lo=0.0), label=rng.integers(0, 10, 8).astype(np.int32), scale=1.0 / 8)
    I["opt"] = dict(n=1000, w=u(9, 1000), g=u(10, 3, 1000), v=u(11, 1000), m=u(12, 1000), v2=u(13, 1000, lo=0.0))
    I["gather"] = dict(x=u(14, 3, 7, 4), index=rng.integers(0, 7, (3, 5, 4)), dout=u(15, 3, 5, 4))
    I["transpose"] = dict(x=u(16, 3, 4, 5), perm=(2, 0, 1))
    I["cast"] = dict(x=u(17, 300) * 100, xi=rng.integers(-1000, 1000, 300).astype(np.int32))
    I["act"] = dict(y=u(18, 500), dy=u(19, 500))
    I["stride"] = dict(inp=u(20, 6, 10), out=u(21, 6, 4), blocks=6, out_blk=4, in_blk=10)
    I["unary"] = dict(x=u(22, 400, lo=0.2, hi=2.0), dy=u(23, 400), scalar=1.7)
    mlab = rng.integers(0, 10, 32).astype(np.int32)
    # the label-metrics kernel asserts exactly one entry > 0.9 per row (metrics_functions.cu:95-102)
    I["metrics"] = dict(logits=u(24, 32, 10, lo=0.0), labels=mlab, dense=np.eye(10, dtype=np.float32)[mlab])
    I["topk"] = dict(x=u(26, 4, 50), k=5, dout=u(27, 4, 5))
    I["replicas"] = dict(x=u(28, 5, 100), n=100, reps=5)
    I["linear"] = dict(x=u(29, 64, 40), w=u(30, 48, 40) / np.sqrt(40).astype(np.float32),
                       # backward: the gradients start non-zero (the reference accumulates with beta = 1)
                       dy=u(33, 64, 48), dx0=u(34, 64, 40) * 0.5, dw0=u(35, 48, 40) * 0.5, db0=u(36, 48) * 0.5)
    I["bmm"] = dict(a=u(31, 10, 10, 5), b=u(32, 10, 5, 10), m=10, n=10, k=5, batch=10,   # test_batch_matmul_kernel.cc
                    do=u(37, 10, 10, 10), da0=u(38, 10, 10, 5) * 0.5, db0=u(39, 10, 5, 10) * 0.5)
    return I


# ------------------------------------------------------------------ backend: reference kernels (GPU)
def run_ref(lib, torch):
    """lib: ctypes.CDLL of libref_kernels.so. Returns {name: numpy array}."""
    I, O = inputs(), {}
    dev = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a)).cuda() if dt is None else \
        torch.from_numpy(np.ascontiguousarray(a)).cuda().to(dt)
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device="cuda")
    P = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    F = ctypes.c_float

    def call(name, *a):
        rc = getattr(lib, name)(*a)
        torch.cuda.synchronize()
        assert rc == 0, f"{name} returned {rc}"

    for tag in ("ln10", "ln600"):
        c = I[tag]
        M, N = c["M"], c["N"]
        x, g, b, dy = dev(c["x"]), dev(c["gamma"]), dev(c["beta"]), dev(c["dy"])
        y, mean, rstd = z(M, N), z(M), z(M)
        call("ref_layer_norm_forward", st, P(x), P(g), P(b), P(y), P(mean), P(rstd), ctypes.c_longlong(M),
             ctypes.c_longlong(N), F(c["eps"]))
        ds, db, sc, bi, dx, dg, dbeta = z(M), z(M), z(M), z(M), z(M, N), z(N), z(N)
        call("ref_layer_norm_backward", st, P(dy), P(x), P(g), P(mean), P(rstd), P(ds), P(db), P(sc), P(bi), P(dx),
             P(dg), P(dbeta), ctypes.c_longlong(M), ctypes.c_longlong(N))
        for k, v in (("y", y), ("mean", mean), ("rstd", rstd), ("dx", dx), ("dgamma", dg), ("dbeta", dbeta)):
            O[f"{tag}_{k}"] = v.cpu().numpy()
    c = I["emb"]
    table, dout = dev(c["table"]), dev(c["dout"])
    for it, idt in (("i32", torch.int32), ("i64", torch.int64)):
        for aggr, an in ((-1, "none"), (0, "sum"), (1, "avg")):
            idx = dev(c["idx1"] if aggr < 0 else c["idx"], idt)
            bag = 1 if aggr < 0 else c["bag"]
            out, dt = z(c["batch"], c["dim"]), z(c["entries"], c["dim"])
            call("ref_embedding_forward", st, P(idx), P(out), P(table), aggr, bag, c["dim"], c["batch"], int(it == "i64"))
            call("ref_embedding_backward", st, P(idx), P(dout), P(dt), aggr, bag, c["dim"], c["batch"], int(it == "i64"))
            O[f"emb_{an}_{it}_out"], O[f"emb_{an}_{it}_dtable"] = out.cpu().numpy(), dt.cpu().numpy()
    c = I["loss"]
    p, y, lab = dev(c["p"]), dev(c["y"]), dev(c["label"])
    g = z(8, 10)
    call("ref_sparse_cce_loss_backward", st, P(g), P(p), P(lab), 8, 10, 1, F(c["scale"]))
    O["loss_sparse_cce"] = g.cpu().numpy()
    for kind, name in ((0, "cce"), (1, "mse"), (2, "identity")):
        g = z(8, 10)
        call("ref_dense_loss_backward", st, kind, P(g), P(p), P(y), ctypes.c_longlong(80), F(c["scale"]))
        O[f"loss_{name}"] = g.cpu().numpy()
    c = I["opt"]
    for name, mom, nest, reps in (("plain", 0.0, 0, 1), ("momentum", 0.9, 0, 3), ("nesterov", 0.9, 1, 1)):
        w, g, v = dev(c["w"]), dev(c["g"]), dev(c["v"])
        call("ref_sgd_update", st, F(0.05), F(mom), nest, F(1e-3), P(g), ctypes.c_size_t(c["n"]), reps, P(w), P(v))
        O[f"sgd_{name}_w"], O[f"sgd_{name}_v"] = w.cpu().numpy(), v.cpu().numpy()
    w, g, m, v = dev(c["w"]), dev(c["g"]), dev(c["m"]), dev(c["v2"])
    call("ref_adam_update", st, F(0.01), F(0.9), F(0.999), F(1e-3), F(1e-8), P(g), ctypes.c_size_t(c["n"]), 2, P(m), P(v), P(w))
    O["adam_w"], O["adam_m"], O["adam_v"] = w.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy()
    c = I["gather"]
    x, dout = dev(c["x"]), dev(c["dout"])
    for it, idt in (("i32", torch.int32), ("i64", torch.int64)):
        idx = dev(c["index"], idt)
        out, din = z(3, 5, 4), z(3, 7, 4)
        args = (ctypes.c_longlong(60), ctypes.c_longlong(4), ctypes.c_longlong(7), ctypes.c_longlong(5), int(it == "i64"))
        call("ref_gather_forward", st, P(x), P(idx), P(out), *args)
        call("ref_gather_backward", st, P(dout), P(idx), P(din), *args)
        O[f"gather_{it}_out"], O[f"gather_{it}_din"] = out.cpu().numpy(), din.cpu().numpy()
    c = I["transpose"]
    x = dev(c["x"])
    shp, perm = c["x"].shape, c["perm"]
    oshp = [shp[p] for p in perm]
    nd = 3
    # legion order: dim 0 innermost; out legion dim i = ff dim nd-1-i, which reads in ff dim perm[nd-1-i]
    ist = [int(np.prod(shp[nd - i:])) for i in range(nd)]
    ost = [int(np.prod(oshp[nd - i:])) for i in range(nd)]
    lperm = [nd - 1 - perm[nd - 1 - i] for i in range(nd)]
    out = z(*oshp)
    A = ctypes.c_int * nd
    call("ref_transpose_kernel", st, P(x), P(out), ctypes.c_longlong(x.numel()), nd, A(*ist), A(*ost), A(*lperm), F(0.0))
    O["transpose_out"] = out.cpu().numpy()
    c = I["cast"]
    x, xi = dev(c["x"]), dev(c["xi"])
    h = z(300, dt=torch.float16)
    call("ref_cast_forward", st, 0, P(x), P(h), ctypes.c_size_t(300))
    O["cast_f2h"] = h.float().cpu().numpy()
    f = z(300)
    call("ref_cast_forward", st, 2, P(xi), P(f), ctypes.c_size_t(300))
    O["cast_i2f"] = f.cpu().numpy()
    i = z(300, dt=torch.int32)
    call("ref_cast_forward", st, 3, P(x), P(i), ctypes.c_size_t(300))
    O["cast_f2i"] = i.cpu().numpy()
    c = I["act"]
    for which, name in ((0, "relu"), (1, "sigmoid")):
        yv = dev(np.maximum(c["y"], 0)) if which == 0 else dev(1 / (1 + np.exp(-c["y"])))
        gr = dev(c["dy"])
        call("ref_activation_backward", st, which, P(gr), P(yv), ctypes.c_size_t(500))
        O[f"act_{name}_bwd"] = gr.cpu().numpy()
    gx = dev(c["y"] * 3)
    call("ref_gelu_forward_inplace", st, P(gx), ctypes.c_size_t(500))
    O["gelu_tanh_fwd"] = gx.cpu().numpy()
    c = I["stride"]
    inp = dev(c["inp"])
    for add in (0, 1):
        out = dev(c["out"])
        call("ref_copy_with_stride", st, P(out), P(inp), c["blocks"], c["out_blk"], c["in_blk"], add)
        O[f"stride_{'add' if add else 'copy'}"] = out.cpu().numpy()
    c = I["unary"]
    x, dy = dev(c["x"]), dev(c["dy"])
    for op in UNARY_OPS:
        y, dx = z(400), z(400)
        call("ref_element_unary_forward", st, REF_UNARY[op], F(c["scalar"]), P(x), P(y), ctypes.c_longlong(400))
        call("ref_element_unary_backward", st, REF_UNARY[op], F(c["scalar"]), P(y), P(dy), P(x), P(dx), ctypes.c_longlong(400))
        O[f"unary_{op}_y"], O[f"unary_{op}_dx"] = y.cpu().numpy(), dx.cpu().numpy()
    c = I["metrics"]
    # (operands are held in variables: a temporary would be freed, and its memory reused, before the launch)
    logits, labels, dense = dev(c["logits"]), dev(c["labels"]), dev(c["dense"])
    perf = z(7, dt=torch.int32)
    call("ref_metrics_sparse", st, 1 | 4 | 8 | 16 | 32, P(logits), P(labels), 32, 10, P(perf))
    O["metrics_sparse"] = perf.cpu().numpy()
    perf = z(7, dt=torch.int32)
    call("ref_metrics_label", st, 1 | 2 | 8 | 16 | 32, P(logits), P(dense), 32, 10, P(perf))
    O["metrics_label"] = perf.cpu().numpy()
    c = I["topk"]
    x = dev(c["x"])
    out, ind, din = z(4, 5), z(4, 5, dt=torch.int32), z(4, 50)
    call("ref_topk_forward", st, P(x), P(out), P(ind), ctypes.c_size_t(4), 50, 5, 1)
    tdout = dev(c["dout"])
    call("ref_topk_backward", st, P(tdout), P(ind), P(din), ctypes.c_size_t(4), 50, 5)
    O["topk_values"], O["topk_indices"], O["topk_din"] = out.cpu().numpy(), ind.cpu().numpy(), din.cpu().numpy()
    c = I["replicas"]
    acc, rx = dev(np.ones(100, np.float32)), dev(c["x"])
    call("ref_replicate_backward", st, P(acc), P(rx), ctypes.c_size_t(100), ctypes.c_size_t(5))
    O["replicate_bwd"] = acc.cpu().numpy()
    red = z(100)
    call("ref_reduction_forward", st, P(rx), P(red), ctypes.c_size_t(100), ctypes.c_size_t(5))
    O["reduction_fwd"] = red.cpu().numpy()
    c = I["linear"]
    for code, name, dt in ((0, "f32", torch.float32), (1, "f16", torch.float16), (2, "bf16", torch.bfloat16)):
        x, w = dev(c["x"], dt), dev(c["w"], dt)
        y = z(64, 48, dt=dt)
        call("ref_linear_forward_gemm", st, P(x), P(w), P(y), 40, 48, 64, code)
        O[f"linear_{name}_y"] = y.float().cpu().numpy()
        dy, dx, dw, db = dev(c["dy"], dt), dev(c["dx0"], dt), dev(c["dw0"], dt), dev(c["db0"], dt)
        ones = torch.ones(64, dtype=dt, device="cuda")
        call("ref_linear_backward_gemm", st, P(x), P(dx), P(dy), P(w), P(dw), P(db), P(ones), 40, 48, 64, code)
        O[f"linear_{name}_dx"], O[f"linear_{name}_dw"], O[f"linear_{name}_db"] = (
            dx.float().cpu().numpy(), dw.float().cpu().numpy(), db.float().cpu().numpy())
    c = I["bmm"]
    o, ba, bb = z(10, 10, 10), dev(c["a"]), dev(c["b"])
    call("ref_batch_matmul_forward", st, P(o), P(ba), P(bb), 10, 10, 5, 10)
    O["bmm_o"] = o.cpu().numpy()
    do, da, db = dev(c["do"]), dev(c["da0"]), dev(c["db0"])
    call("ref_batch_matmul_backward", st, P(do), P(ba), P(da), P(bb), P(db), 10, 10, 5, 10)
    O["bmm_da"], O["bmm_db"] = da.cpu().numpy(), db.cpu().numpy()
    return O


# ------------------------------------------------------------------ backend: CPU oracle
def run_oracle(o):
    I, O = inputs(), {}
    z = lambda *s, dt=np.float32: np.zeros(s, dt)
    for tag in ("ln10", "ln600"):
        c = I[tag]
        M, N = c["M"], c["N"]
        y, mean, rstd = z(M, N), z(M), z(M)
        o.layer_norm_forward(c["x"], y, c["gamma"], c["beta"], mean, rstd, M, N, c["eps"])
        dx, dg, db = z(M, N), z(N), z(N)
        o.layer_norm_backward(c["dy"], c["x"], dx, c["gamma"], dg, db, mean, rstd, M, N, 0)
        for k, v in (("y", y), ("mean", mean), ("rstd", rstd), ("dx", dx), ("dgamma", dg), ("dbeta", db)):
            O[f"{tag}_{k}"] = v
    c = I["emb"]
    for it in ("i32", "i64"):   # the oracle takes int64 indices; the index width does not change values
        for aggr, an in ((-1, "none"), (0, "sum"), (1, "avg")):
            idx = np.ascontiguousarray((c["idx1"] if aggr < 0 else c["idx"]).astype(np.int64))
            bag = 1 if aggr < 0 else c["bag"]
            out, dt = z(c["batch"], c["dim"]), z(c["entries"], c["dim"])
            o.embedding_forward(idx, out, c["table"], aggr, bag, c["dim"], c["batch"])
            o.embedding_backward(idx, c["dout"], dt, aggr, bag, c["dim"], c["batch"])
            O[f"emb_{an}_{it}_out"], O[f"emb_{an}_{it}_dtable"] = out, dt
    c = I["loss"]
    g = z(8, 10)
    o.sparse_cce_loss_backward(g, c["p"], c["label"], 8, 10, 1, c["scale"])
    O["loss_sparse_cce"] = g
    for name in ("cce", "mse"):
        g = z(8, 10)
        o.diff_loss_backward(g, c["p"], c["y"], 80, c["scale"])
        O[f"loss_{name}"] = g
    O["loss_identity"] = np.full((8, 10), c["scale"], np.float32)
    c = I["opt"]
    for name, mom, nest, reps in (("plain", 0.0, 0, 1), ("momentum", 0.9, 0, 3), ("nesterov", 0.9, 1, 1)):
        w, g, v = c["w"].copy(), c["g"].copy(), c["v"].copy()
        o.sgd_update(0.05, mom, nest, 1e-3, g, c["n"], reps, w, v)
        O[f"sgd_{name}_w"], O[f"sgd_{name}_v"] = w, v
    w, g, m, v = c["w"].copy(), c["g"].copy(), c["m"].copy(), c["v2"].copy()
    o.adam_update(0.01, 0.9, 0.999, 1e-3, 1e-8, g, c["n"], 2, m, v, w)
    O["adam_w"], O["adam_m"], O["adam_v"] = w, m, v
    c = I["gather"]
    for it in ("i32", "i64"):
        idx = np.ascontiguousarray(c["index"].astype(np.int64))
        out, din = z(3, 5, 4), z(3, 7, 4)
        o.gather_forward(c["x"], idx, out, 60, 4, 7, 5)
        o.gather_backward(c["dout"], idx, din, 60, 4, 7, 5)
        O[f"gather_{it}_out"], O[f"gather_{it}_din"] = out, din
    c = I["transpose"]
    out = z(*[c["x"].shape[p] for p in c["perm"]])
    o.transpose(c["x"], out, 3, list(c["x"].shape), list(c["perm"]), 0)
    O["transpose_out"] = out
    c = I["cast"]
    O["cast_f2h"] = c["x"].astype(np.float16).astype(np.float32)
    O["cast_i2f"] = c["xi"].astype(np.float32)
    O["cast_f2i"] = np.trunc(c["x"]).astype(np.int32)
    c = I["act"]
    # Linear's activation gradients (linear_backward's in-place step) through the oracle
    for act, name, yv in ((0, "relu", np.maximum(c["y"], 0)), (1, "sigmoid", (1 / (1 + np.exp(-c["y"]))).astype(np.float32))):
        dy = c["dy"].copy().reshape(1, 500)
        x1, w1 = z(1, 1), z(500, 1)
        o.linear_backward(x1, None, np.ascontiguousarray(yv.reshape(1, 500)), dy, w1, z(500, 1), None, 1, 500, 1, act, 0.0, None)
        O[f"act_{name}_bwd"] = dy.reshape(500)
    yg = z(1, 500)
    o.linear_forward(np.ones((1, 1), np.float32), np.ascontiguousarray((c["y"] * 3).reshape(500, 1)), None, yg, 1, 500, 1, 3)
    O["gelu_tanh_fwd"] = yg.reshape(500)
    c = I["stride"]
    for add in (0, 1):
        out = c["out"].copy()
        o.copy_with_stride(out, c["inp"], c["blocks"], c["out_blk"], c["in_blk"], add)
        O[f"stride_{'add' if add else 'copy'}"] = out
    c = I["unary"]
    for op in UNARY_OPS:
        y, dx = z(400), z(400)
        o.element_unary_forward(FF_UNARY[op], c["scalar"], c["x"], y, 400)
        o.element_unary_backward(FF_UNARY[op], c["scalar"], c["x"], dx, y, c["dy"], 400)
        O[f"unary_{op}_y"], O[f"unary_{op}_dx"] = y, dx
    c = I["metrics"]
    for name, flags, fn, lab in (("sparse", 1 | 4 | 8 | 16 | 32, o.metrics_sparse, c["labels"]),
                                 ("label", 1 | 2 | 8 | 16 | 32, o.metrics_label, c["dense"])):
        acc = np.zeros(7, np.float64)
        fn(flags, c["logits"], lab, 32, 10, acc)
        raw = np.zeros(7, np.int32)
        raw[:2] = acc[:2].astype(np.int32)
        raw[2:] = acc[2:].astype(np.float32).view(np.int32)
        O[f"metrics_{name}"] = raw
    c = I["topk"]
    out, ind, din = z(4, 5), z(4, 5, dt=np.int32), z(4, 50)
    o.topk_forward(c["x"], out, ind, 4, 50, 5)
    o.topk_backward(c["dout"], ind, din, 4, 50, 5)
    O["topk_values"], O["topk_indices"], O["topk_din"] = out, ind, din
    c = I["replicas"]
    acc = np.ones(100, np.float32)
    o.sum_replicas(acc, c["x"], 100, 5, 1)
    O["replicate_bwd"] = acc
    red = z(100)
    o.sum_replicas(red, c["x"], 100, 5, 0)
    O["reduction_fwd"] = red
    c = I["linear"]
    for name, dt in (("f32", np.float32), ("f16", np.float16), ("bf16", None)):
        x, w = c["x"], c["w"]
        if dt is np.float16:
            x, w = x.astype(np.float16).astype(np.float32), w.astype(np.float16).astype(np.float32)
        elif dt is None:
            x, w = _bf16(x), _bf16(w)
        y = z(64, 48)
        o.linear_forward(np.ascontiguousarray(x), np.ascontiguousarray(w), None, y, 40, 48, 64, -1)
        O[f"linear_{name}_y"] = y
        rnd = (lambda a: a.astype(np.float16).astype(np.float32)) if dt is np.float16 else _bf16 if dt is None else (lambda a: a)
        dy, dx, dw, db = (np.ascontiguousarray(rnd(c[k])).copy() for k in ("dy", "dx0", "dw0", "db0"))
        o.linear_backward(np.ascontiguousarray(x), dx, y, dy, np.ascontiguousarray(w), dw, db, 40, 48, 64, -1, 0.0, None)
        O[f"linear_{name}_dx"], O[f"linear_{name}_dw"], O[f"linear_{name}_db"] = dx, dw, db
    c = I["bmm"]
    out = z(10, 10, 10)
    o.batch_matmul_forward(out, c["a"], c["b"], 10, 10, 5, 10, -1, -1, -1)
    O["bmm_o"] = out
    da, db = c["da0"].copy(), c["db0"].copy()
    o.batch_matmul_backward(c["do"], c["a"], da, c["b"], db, 10, 10, 5, 10)
    O["bmm_da"], O["bmm_db"] = da, db
    return O


def _bf16(a):
    """Round fp32 to bf16 (nearest even) and back."""
    b = np.ascontiguousarray(a, np.float32).view(np.uint32)
    r = ((b + 0x7FFF + ((b >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
    return r.view(np.float32)


# ------------------------------------------------------------------ backend: libffk.so (GPU)
def run_ffk(K, torch):
    I, O = inputs(), {}
    dev = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a)).cuda() if dt is None else \
        torch.from_numpy(np.ascontiguousarray(a)).cuda().to(dt)
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device="cuda")
    for tag in ("ln10", "ln600"):
        c = I[tag]
        M, N = c["M"], c["N"]
        x, g, b, dy = dev(c["x"]), dev(c["gamma"]), dev(c["beta"]), dev(c["dy"])
        ln = K.LayerNorm.init_kernel(True, M, N, c["eps"])
        y, dx, dg, db = z(M, N), z(M, N), z(N), z(N)
        K.LayerNorm.forward_kernel(None, ln, x, y, g, b)
        K.LayerNorm.backward_kernel(None, ln, dy, x, dx, g, dg, db)
        for k, v in (("y", y), ("mean", ln.mean[:M]), ("rstd", ln.rstd[:M]), ("dx", dx), ("dgamma", dg), ("dbeta", db)):
            O[f"{tag}_{k}"] = v.float().cpu().numpy()
    c = I["emb"]
    table, dout = dev(c["table"]), dev(c["dout"])
    for it, idt in (("i32", torch.int32), ("i64", torch.int64)):
        for aggr, an in ((K.AGGR_NONE, "none"), (K.AGGR_SUM, "sum"), (K.AGGR_AVG, "avg")):
            none = an == "none"
            idx = dev(c["idx1"] if none else c["idx"], idt)
            bag = 1 if none else c["bag"]
            out, dt = z(c["batch"], c["dim"]), z(c["entries"], c["dim"])
            K.Embedding.forward_kernel(None, idx, out, table, aggr, bag, c["dim"], c["batch"])
            K.Embedding.backward_kernel(None, idx, dout, dt, aggr, bag, c["dim"], c["batch"])
            O[f"emb_{an}_{it}_out"], O[f"emb_{an}_{it}_dtable"] = out.cpu().numpy(), dt.cpu().numpy()
    c = I["loss"]
    p, y, lab = dev(c["p"]), dev(c["y"]), dev(c["label"])
    g = z(8, 10)
    K.sparse_categorical_crossentropy_loss_backward_kernel(None, g, p, lab, 8, 10, 1, c["scale"])
    O["loss_sparse_cce"] = g.cpu().numpy()
    g = z(8, 10)
    K.categorical_crossentropy_loss_backward_kernel(None, g, p, y, c["scale"])
    O["loss_cce"] = g.cpu().numpy()
    g = z(8, 10)
    K.mean_squared_error_avg_loss_backward_kernel(None, g, p, y, c["scale"])
    O["loss_mse"] = g.cpu().numpy()
    g = z(8, 10)
    K.identity_loss_backward_kernel(None, g, p, c["scale"])
    O["loss_identity"] = g.cpu().numpy()
    c = I["opt"]
    for name, mom, nest, reps in (("plain", 0.0, 0, 1), ("momentum", 0.9, 0, 3), ("nesterov", 0.9, 1, 1)):
        w, g, v = dev(c["w"]), dev(c["g"]), dev(c["v"])
        K.sgd_ps_update_task_gpu(None, 0.05, mom, nest, 1e-3, g, c["n"], reps, w, v)
        O[f"sgd_{name}_w"], O[f"sgd_{name}_v"] = w.cpu().numpy(), v.cpu().numpy()
    w, g, m, v = dev(c["w"]), dev(c["g"]), dev(c["m"]), dev(c["v2"])
    K.adam_ps_update_task_gpu(None, 0.01, 0.9, 0.999, 1e-3, 1e-8, g, c["n"], 2, m, v, w)
    O["adam_w"], O["adam_m"], O["adam_v"] = w.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy()
    c = I["gather"]
    x, dout = dev(c["x"]), dev(c["dout"])
    for it, idt in (("i32", torch.int32), ("i64", torch.int64)):
        idx = dev(c["index"], idt)
        out, din = z(3, 5, 4), z(3, 7, 4)
        K.Gather.forward_kernel(None, 1, x, idx, out)
        K.Gather.backward_kernel(None, 1, dout, idx, din)
        O[f"gather_{it}_out"], O[f"gather_{it}_din"] = out.cpu().numpy(), din.cpu().numpy()
    c = I["transpose"]
    x = dev(c["x"])
    out = z(*[c["x"].shape[p] for p in c["perm"]])
    K.Transpose.forward_kernel(None, K.Transpose.init_kernel(3, c["perm"]), x, out)
    O["transpose_out"] = out.cpu().numpy()
    c = I["cast"]
    x, xi = dev(c["x"]), dev(c["xi"])
    h, f, i = z(300, dt=torch.float16), z(300), z(300, dt=torch.int32)
    K.Cast.forward_kernel(None, x, h)
    K.Cast.forward_kernel(None, xi, f)
    K.Cast.forward_kernel(None, x, i)
    O["cast_f2h"], O["cast_i2f"], O["cast_f2i"] = h.float().cpu().numpy(), f.cpu().numpy(), i.cpu().numpy()
    c = I["act"]
    for act, name, yv in ((K.ACT_RELU, "relu", np.maximum(c["y"], 0)),
                          (K.ACT_SIGMOID, "sigmoid", (1 / (1 + np.exp(-c["y"]))).astype(np.float32))):
        lin = K.Linear.init_kernel(act, 0.0, False, K.GEMM_SIMT)
        dy = dev(c["dy"]).view(1, 500).contiguous()
        K.Linear.backward_kernel(None, lin, z(1, 1), None, dev(yv).view(1, 500).contiguous(), dy, z(500, 1), z(500, 1), None, 1, 500, 1)
        O[f"act_{name}_bwd"] = dy.cpu().numpy().reshape(500)
    lin = K.Linear.init_kernel(K.ACT_GELU, 0.0, False, K.GEMM_SIMT)
    yg = z(1, 500)
    K.Linear.forward_kernel(None, lin, torch.ones(1, 1, device="cuda"), yg, dev(c["y"] * 3).view(500, 1).contiguous(), None, 1, 500, 1)
    O["gelu_tanh_fwd"] = yg.cpu().numpy().reshape(500)
    c = I["stride"]
    inp = dev(c["inp"])
    for add in (0, 1):
        out = dev(c["out"])
        K._call("ffk_add_with_stride" if add else "ffk_copy_with_stride", K._stream(None), out.data_ptr(),
                inp.data_ptr(), c["blocks"], c["out_blk"], c["in_blk"], K.dtype_code(out))
        O[f"stride_{'add' if add else 'copy'}"] = out.cpu().numpy()
    c = I["unary"]
    x, dy = dev(c["x"]), dev(c["dy"])
    for op in UNARY_OPS:
        y, dx = z(400), z(400)
        K.ElementUnary.forward_kernel(None, FF_UNARY[op], c["scalar"], x, y)
        K.ElementUnary.backward_kernel(None, FF_UNARY[op], c["scalar"], x, dx, y, dy)
        O[f"unary_{op}_y"], O[f"unary_{op}_dx"] = y.cpu().numpy(), dx.cpu().numpy()
    c = I["metrics"]
    perf = K.PerfMetrics()
    K.update_metrics_sparse_label_kernel(None, 1 | 4 | 8 | 16 | 32, dev(c["logits"]), dev(c["labels"]), 32, 10, perf)
    O["metrics_sparse"] = perf.buf.cpu().numpy()
    perf = K.PerfMetrics()
    K.update_metrics_label_kernel(None, 1 | 2 | 8 | 16 | 32, dev(c["logits"]), dev(c["dense"]), 32, 10, perf)
    O["metrics_label"] = perf.buf.cpu().numpy()
    c = I["topk"]
    x = dev(c["x"])
    out, ind, din = z(4, 5), z(4, 5, dt=torch.int32), z(4, 50)
    K.TopK.forward_kernel(None, x, out, ind, 4, 50, 5, True)
    K.TopK.backward_kernel(None, dev(c["dout"]), ind, din, 4, 50, 5)
    O["topk_values"], O["topk_indices"], O["topk_din"] = out.cpu().numpy(), ind.cpu().numpy(), din.cpu().numpy()
    c = I["replicas"]
    acc = dev(np.ones(100, np.float32))
    K.Replicate.backward_kernel(None, acc, dev(c["x"]), 5)
    O["replicate_bwd"] = acc.cpu().numpy()
    red = z(100)
    K.Reduction.forward_kernel(None, dev(c["x"]), red, 5)
    O["reduction_fwd"] = red.cpu().numpy()
    c = I["linear"]
    for name, dt in (("f32", torch.float32), ("f16", torch.float16), ("bf16", torch.bfloat16)):
        lin = K.Linear.init_kernel(K.ACT_NONE, 0.0, False, K.GEMM_AUTO)
        x, w = dev(c["x"], dt), dev(c["w"], dt)
        y = z(64, 48, dt=dt)
        K.Linear.forward_kernel(None, lin, x, y, w, None, 40, 48, 64)
        O[f"linear_{name}_y"] = y.float().cpu().numpy()
        linb = K.Linear.init_kernel(K.ACT_NONE, 0.0, True, K.GEMM_AUTO, out_dim=48)
        dy, dx, dw, db = dev(c["dy"], dt), dev(c["dx0"], dt), dev(c["dw0"], dt), dev(c["db0"], dt)
        K.Linear.backward_kernel(None, linb, x, dx, y, dy, w, dw, db, 40, 48, 64)
        O[f"linear_{name}_dx"], O[f"linear_{name}_dw"], O[f"linear_{name}_db"] = (
            dx.float().cpu().numpy(), dw.float().cpu().numpy(), db.float().cpu().numpy())
    c = I["bmm"]
    o = z(10, 10, 10)
    K.BatchMatmul.forward_kernel(None, o, dev(c["a"]), dev(c["b"]), 10, 10, 5, 10)
    O["bmm_o"] = o.cpu().numpy()
    da, db = dev(c["da0"]), dev(c["db0"])
    K.BatchMatmul.backward_kernel(None, o, dev(c["do"]), dev(c["a"]), da, dev(c["b"]), db, 10, 10, 5, 10)
    O["bmm_da"], O["bmm_db"] = da.cpu().numpy(), db.cpu().numpy()
    torch.cuda.synchronize()
    return O


# ------------------------------------------------------------------ comparison rule
EXACT = ("emb_none_", "gather_i32_out", "gather_i64_out", "transpose_out", "cast_", "stride_copy", "topk_indices",
         "topk_values", "loss_identity")
# fp16 with CUBLAS_COMPUTE_16F (the reference's choice) accumulates in half: compare at half resolution
RTOL = {"linear_f16_y": 2e-2, "linear_bf16_y": 2e-2, "unary_POW_y": 1e-5, "unary_POW_dx": 1e-5}
RTOL.update({f"linear_{t}_{g}": 2e-2 for t in ("f16", "bf16") for g in ("dx", "dw", "db")})


def compare(name, got, ref):
    """None when `got` agrees with the golden `ref` under the rule for `name`, else a message."""
    got, ref = np.asarray(got), np.asarray(ref)
    if got.shape != ref.shape:
        got = got.reshape(ref.shape)
    if name.startswith("metrics_"):
        gi, ri = got.view(np.int32), ref.view(np.int32)
        if not (gi[:2] == ri[:2]).all():
            return f"{name}: counters {gi[:2]} vs {ri[:2]}"
        gf, rf = gi[2:].view(np.float32), ri[2:].view(np.float32)
        return None if np.allclose(gf, rf, rtol=1e-5, atol=1e-6) else f"{name}: {gf} vs {rf}"
    if any(name.startswith(p) or name == p for p in EXACT):
        return None if np.array_equal(got, ref) else f"{name}: not bit-identical ({(got != ref).sum()} elements)"
    rtol = RTOL.get(name, 1e-4)
    atol = rtol * float(np.mean(np.abs(ref))) + 1e-30
    bad = np.abs(got.astype(np.float64) - ref.astype(np.float64)) > atol + rtol * np.abs(ref)
    return None if not bad.any() else f"{name}: {bad.sum()} of {bad.size} elements outside rtol {rtol}"
