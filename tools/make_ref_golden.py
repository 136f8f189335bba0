#!/usr/bin/env python
"""Runs the reference's own CUDA kernels (oracle/_ref/libref_kernels.so, built here from the
reference checkout by `make -C oracle/ref_kernels`) on the seeded cases of tests/ref_cases.py and
writes gpurun_out/ref_kernels.npz. Run on the GPU box:

    gpurun -- python tools/make_ref_golden.py

then copy the file to tests/golden/ref_kernels.npz and commit it (`--merge` keeps the arrays that
are already committed and only adds new cases). /root/reference is not read at
run time: the .so travels with the snapshot."""
import ctypes
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tests"))
import ref_cases  # noqa: E402


def main():
    lib = ctypes.CDLL(str(ref_cases.REF_SO))
    out = ref_cases.run_ref(lib, torch)
    if "--merge" in sys.argv:
        # cases added later: arrays already committed stay byte for byte, only new names are added
        old = dict(np.load(ROOT / "tests" / "golden" / "ref_kernels.npz"))
        added = sorted(set(out) - set(old))
        out = {**out, **old}
        print(f"merge: kept {len(old)} committed arrays, added {added}")
    dest = ROOT / "gpurun_out" / "ref_kernels.npz"
    dest.parent.mkdir(exist_ok=True)
    np.savez_compressed(dest, **out)
    print(f"wrote {dest}: {len(out)} arrays, {sum(v.nbytes for v in out.values())} bytes; "
          f"GPU {torch.cuda.get_device_name(0)}, CUDA {torch.version.cuda}")


if __name__ == "__main__":
    main()
