"""Turns the raw dumps of oracle/cudnn_ref/mha_golden.cu (run on the GPU box, see that file)
into tests/golden/mha_<case>.npz, and checks a numpy statement of the math against them.

    gpurun -- 'oracle/_build/mha_golden gpurun_out/mha_<case> N H qSize kSize vSize qProj kProj vProj oProj qoSeq kvSeq'
    python tools/make_mha_golden.py gpurun_out/mha_fixture gpurun_out/mha_ragged gpurun_out/mha_cross
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]


def mha_numpy(c, q, k, v, w):
    """cuDNN MHA as configured by the reference: no biases, scale 1.0, full window."""
    N, H = c["N"], c["H"]
    q = q.reshape(N, c["qoSeq"], c["qSize"]).astype(np.float64)
    k = k.reshape(N, c["kvSeq"], c["kSize"]).astype(np.float64)
    v = v.reshape(N, c["kvSeq"], c["vSize"]).astype(np.float64)
    w = w.astype(np.float64)
    o = 0
    wq = w[o:o + c["qSize"] * H * c["qProj"]].reshape(c["qSize"], H, c["qProj"]); o += wq.size
    wk = w[o:o + c["kSize"] * H * c["kProj"]].reshape(c["kSize"], H, c["kProj"]); o += wk.size
    wv = w[o:o + c["vSize"] * H * c["vProj"]].reshape(c["vSize"], H, c["vProj"]); o += wv.size
    wo = w[o:o + H * c["vProj"] * c["oProj"]].reshape(H, c["vProj"], c["oProj"])
    qp = np.einsum("nsi,ihp->nhsp", q, wq)
    kp = np.einsum("nti,ihp->nhtp", k, wk)
    vp = np.einsum("nti,ihp->nhtp", v, wv)
    s = np.einsum("nhsp,nhtp->nhst", qp, kp)  # smScalar = 1.0
    s = s - s.max(-1, keepdims=True)
    p = np.exp(s)
    p /= p.sum(-1, keepdims=True)
    oh = np.einsum("nhst,nhtp->nshp", p, vp)
    return np.einsum("nshp,hpo->nso", oh, wo)


def main(dirs):
    out_dir = ROOT / "tests" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    for d in dirs:
        d = Path(d)
        layout = json.loads((d / "layout.json").read_text())
        arrs = {n: np.fromfile(d / f"{n}.bin", dtype=np.float32)
                for n in ("q", "k", "v", "w", "dout", "out", "dq", "dk", "dv", "dw")}
        c = layout["config"]
        ref = mha_numpy(c, arrs["q"], arrs["k"], arrs["v"], arrs["w"]).reshape(-1)
        err = np.abs(ref - arrs["out"]).max() / np.abs(arrs["out"]).max()
        print(f"{d.name}: numpy restatement vs cuDNN forward: max rel err {err:.2e}")
        name = d.name.replace("mha_", "")
        np.savez_compressed(out_dir / f"mha_{name}.npz", layout=json.dumps(layout), **arrs)
        print("  wrote", out_dir / f"mha_{name}.npz")


if __name__ == "__main__":
    main(sys.argv[1:])
