"""Peer-memory all-reduce (ffk_p2p_*, include/ffk_parallel.h): world 1 on any GPU box, world 2, 4
and 8 when the box has that many GPUs (one process per GPU, handles exchanged over gloo). All
three protocols (push <= 64 KB, one-shot <= 512 KB, two-shot) and the background kernel must give
results bit-identical to the fp32 rank-order sum; a rank that never arrives must turn into NaN
results and an error status on every rank, not into a partial sum."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
WORKER = Path(__file__).resolve().parent / "p2p_worker.py"


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world, worker=WORKER, extra_env=None, args=()):
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), **(extra_env or {}))
        procs.append(subprocess.Popen([sys.executable, str(worker), *args], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            pytest.fail("p2p worker timed out")
        outs.append(out)
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {rank} failed:\n{out[-2000:]}"


def test_p2p_all_reduce_single_rank():
    _run(1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_p2p_all_reduce_multi_rank(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs on one node")
    _run(world)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one node")
def test_p2p_missing_rank_poisons_the_group():
    """Rank 1 skips one collective. Rank 0's wait expires (FFK_P2P_TIMEOUT_MS = 400): it must
    publish NaN, not a partial sum; when rank 1 calls its next collective it must find the group
    poisoned and get NaN and an error status as well."""
    _run(2, extra_env={"FFK_P2P_TIMEOUT_MS": "400"}, args=("timeout",))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_data_parallel_encoder_peer_sums_match_nccl(world):
    """A toy encoder trained data-parallel for three steps twice: gradient sums over peer memory
    issued under the backward pass (PeerGradSync) against NCCL all-reduces after backward. The
    weights must agree to one bf16 rounding, and be identical on all ranks."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs on one node")
    _run(world, worker=Path(__file__).resolve().parent / "dp_worker.py")
