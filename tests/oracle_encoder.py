"""The transformer-encoder training step of BASELINE configs[2] composed from the CPU oracle's
functions (oracle/cpu_ref.cc) — test infrastructure: the referee of the encoder parity test and
the CPU leg of bench.py (`cpu_baseline`, `--impl reference`). One layer is what lib/models builds
(lib/models/src/models/transformer/transformer.cc:32-65):

    MHA(x,x,x) -> add -> LayerNorm -> dense(dim_ff, ReLU) -> dropout -> dense(d) -> dropout -> add -> LayerNorm

followed by the MSE-average loss gradient (scale 2/volume), backward in reverse order (the
attention input gradient is dq + dk + dv), and SGD with weight decay. All arrays are fp32."""
import numpy as np

EW_ADD = 28  # OperatorType::EW_ADD (op-attrs/operator_type.enum.toml order, as kernels.py uses it)


class OracleEncoder:
    def __init__(self, o, weights, batch, seq, d_model, heads, head_dim, dim_ff, eps=1e-5, rate=0.0):
        """weights: per layer a dict of fp32 arrays: mha (packed), g1, b1, w1, c1, w2, c2, g2, b2."""
        self.o, self.W = o, weights
        self.B, self.S, self.D, self.H, self.hd, self.FF = batch, seq, d_model, heads, head_dim, dim_ff
        self.eps, self.rate = eps, rate
        self.dims = [batch, heads, d_model, d_model, d_model, head_dim, head_dim, head_dim, d_model, seq, seq]

    @staticmethod
    def random_weights(layers, d_model, heads, head_dim, dim_ff, seed=0):
        rng = np.random.default_rng(seed)
        u = lambda n, s: (rng.uniform(-1, 1, n) * s).astype(np.float32)
        out = []
        nw = 3 * d_model * heads * head_dim + heads * head_dim * d_model
        for _ in range(layers):
            out.append(dict(mha=u(nw, d_model ** -0.5), g1=np.ones(d_model, np.float32), b1=np.zeros(d_model, np.float32),
                            w1=u((dim_ff, d_model), (6 / (d_model + dim_ff)) ** 0.5), c1=np.zeros(dim_ff, np.float32),
                            w2=u((d_model, dim_ff), (6 / (d_model + dim_ff)) ** 0.5), c2=np.zeros(d_model, np.float32),
                            g2=np.ones(d_model, np.float32), b2=np.zeros(d_model, np.float32)))
        return out

    def step(self, x, target, lr, wd, masks=None, global_volume=None):
        """One training step in place on self.W; returns the model output of the forward pass.
        masks: per layer (bits1, bits2) uint32 words of the two dropout layers (rate > 0)."""
        o, B, S, D, FF = self.o, self.B, self.S, self.D, self.FF
        M = B * S
        z = lambda *s: np.zeros(s, np.float32)
        shp = [M, D]
        saved = []
        cur = np.ascontiguousarray(x, np.float32).reshape(M, D)
        for li, w in enumerate(self.W):
            att = z(M, D)
            o.mha_forward(self.dims, cur, cur, cur, w["mha"], att)
            a1 = z(M, D)
            o.element_binary_forward(EW_ADD, cur, att, a1, 2, shp, shp, shp)
            n1, mean1, rstd1 = z(M, D), z(M), z(M)
            o.layer_norm_forward(a1, n1, w["g1"], w["b1"], mean1, rstd1, M, D, self.eps)
            h = z(M, FF)
            o.linear_forward(n1, w["w1"], w["c1"], h, D, FF, M, 0)          # ReLU
            hd = h
            if self.rate > 0:
                hd = z(M, FF)
                o.dropout_apply(h, hd, masks[li][0], h.size, self.rate)
            f = z(M, D)
            o.linear_forward(hd, w["w2"], w["c2"], f, FF, D, M, -1)
            fd = f
            if self.rate > 0:
                fd = z(M, D)
                o.dropout_apply(f, fd, masks[li][1], f.size, self.rate)
            a2 = z(M, D)
            o.element_binary_forward(EW_ADD, n1, fd, a2, 2, shp, shp, shp)
            n2, mean2, rstd2 = z(M, D), z(M), z(M)
            o.layer_norm_forward(a2, n2, w["g2"], w["b2"], mean2, rstd2, M, D, self.eps)
            saved.append((cur, a1, n1, mean1, rstd1, h, hd, f, a2, mean2, rstd2))
            cur = n2
        out = cur
        g = z(M, D)
        scale = 2.0 / float(global_volume or out.size)
        o.diff_loss_backward(g, out, np.ascontiguousarray(target, np.float32).reshape(M, D), out.size, scale)
        for li in reversed(range(len(self.W))):
            w = self.W[li]
            x0, a1, n1, mean1, rstd1, h, hd, f, a2, mean2, rstd2 = saved[li]
            dg2, db2, da2 = z(D), z(D), z(M, D)
            o.layer_norm_backward(g, a2, da2, w["g2"], dg2, db2, mean2, rstd2, M, D, 0)
            dn1, dfd = z(M, D), z(M, D)
            o.element_binary_backward(EW_ADD, da2, n1, f, dn1, dfd, 2, shp, shp, shp)
            df = dfd
            if self.rate > 0:
                df = z(M, D)
                o.dropout_apply(dfd, df, masks[li][1], df.size, self.rate)
            dhd, dw2, dc2 = z(M, FF), z(D, FF), z(D)
            o.linear_backward(hd, dhd, f, df, w["w2"], dw2, dc2, FF, D, M, -1, 0.0, None)
            dh = dhd
            if self.rate > 0:
                dh = z(M, FF)
                o.dropout_apply(dhd, dh, masks[li][0], dh.size, self.rate)
            dw1, dc1 = z(FF, D), z(FF)
            o.linear_backward(n1, dn1, h, dh, w["w1"], dw1, dc1, D, FF, M, 0, 0.0, None)
            dg1, db1, da1 = z(D), z(D), z(M, D)
            o.layer_norm_backward(dn1, a1, da1, w["g1"], dg1, db1, mean1, rstd1, M, D, 0)
            dx, datt = z(M, D), z(M, D)
            o.element_binary_backward(EW_ADD, da1, x0, x0, dx, datt, 2, shp, shp, shp)
            dq, dk, dv, dmha = z(M, D), z(M, D), z(M, D), np.zeros_like(w["mha"])
            o.mha_backward(self.dims, x0, dq, x0, dk, x0, dv, w["mha"], dmha, datt)
            g = dx + dq + dk + dv
            for name, grad in (("mha", dmha), ("g1", dg1), ("b1", db1), ("w1", dw1), ("c1", dc1), ("w2", dw2),
                               ("c2", dc2), ("g2", dg2), ("b2", db2)):
                o.sgd_update(lr, 0.0, 0, wd, grad.reshape(-1), grad.size, 1, w[name].reshape(-1), None)
        return out.reshape(B, S, D)
