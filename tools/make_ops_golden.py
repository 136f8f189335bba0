"""Packs the raw dumps of oracle/cudnn_ref/ops_golden.cu (run on the GPU box) into
tests/golden/cudnn_ops.npz.

    gpurun -- 'mkdir -p gpurun_out/ops_golden && oracle/_build/ops_golden gpurun_out/ops_golden'
    python tools/make_ops_golden.py gpurun_out/ops_golden

A second argument names the file (default cudnn_ops.npz), e.g. for oracle/cudnn_ref/conv_golden.cu:

    gpurun -- 'mkdir -p gpurun_out/conv_golden && oracle/_build/conv_golden gpurun_out/conv_golden'
    python tools/make_ops_golden.py gpurun_out/conv_golden cudnn_conv.npz
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]


def main(d, name="cudnn_ops.npz"):
    d = Path(d)
    meta = json.loads((d / "meta.json").read_text())
    arrs = {name: np.fromfile(d / f"{name}.bin", dtype=np.float32).reshape(shape) for name, shape in meta.items()}
    out = ROOT / "tests" / "golden" / name
    np.savez_compressed(out, **arrs)
    print("wrote", out, "with", len(arrs), "arrays")


if __name__ == "__main__":
    main(*sys.argv[1:3])
