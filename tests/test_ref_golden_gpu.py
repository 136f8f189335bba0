"""libffk.so against tests/golden/ref_kernels.npz (outputs of the reference's own CUDA kernels and
cuBLAS calls on a B200, see test_ref_golden_cpu.py), through the C ABI. Where
oracle/_ref/libref_kernels.so travelled to this box the reference kernels are also run LIVE and must
reproduce the committed file. Deviations (ref_cases.DEVIATIONS) are judged against the oracle."""
import ctypes
from pathlib import Path

import numpy as np
import pytest
import torch

import ref_cases

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden" / "ref_kernels.npz"


@pytest.fixture(scope="module")
def golden():
    if not GOLDEN.exists():
        pytest.skip("tests/golden/ref_kernels.npz has not been generated yet (tools/make_ref_golden.py)")
    return dict(np.load(GOLDEN))


def test_kernels_match_the_reference_kernels(golden, K, oracle):
    got = ref_cases.run_ffk(K, torch)
    want_dev = ref_cases.run_oracle(oracle)
    failures = []
    for name in sorted(golden):
        ref = want_dev[name] if name in ref_cases.DEVIATIONS else golden[name]
        msg = ref_cases.compare(name, got[name], ref)
        if msg is not None:
            failures.append(msg)
    assert not failures, "\n".join(failures)


def test_live_reference_kernels_reproduce_the_golden_file(golden):
    if not ref_cases.REF_SO.exists():
        pytest.skip("oracle/_ref/libref_kernels.so was not built (needs the reference checkout at build time)")
    live = ref_cases.run_ref(ctypes.CDLL(str(ref_cases.REF_SO)), torch)
    failures = [m for m in (ref_cases.compare(n, live[n], golden[n]) for n in sorted(golden)) if m]
    assert not failures, "\n".join(failures)
