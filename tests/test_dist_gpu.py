"""BASELINE config 4 in miniature on real GPUs: successive halving over a live population sharded over 2 ranks
(torchrun, one process per GPU), evaluation scores all-gathered over NCCL, members migrated between the GPUs (one NCCL
send / recv of the member's state each), identical pruning on every rank and equal to the single-process run. Needs >= 2 GPUs (skipped otherwise)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_successive_halving_matches_single_process(tmp_path):
    worker = os.path.join(HERE, "dist_gpu_worker.py")
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    single = subprocess.run([sys.executable, worker, str(tmp_path)], env=env, capture_output=True, text=True, timeout=600)
    assert single.returncode == 0, single.stdout[-2000:] + single.stderr[-2000:]
    multi = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                            "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), worker, str(tmp_path)],
                           env=env, capture_output=True, text=True, timeout=600)
    assert multi.returncode == 0, multi.stdout[-2000:] + multi.stderr[-2000:]
    s = json.load(open(tmp_path / "w1r0.json"))
    r0, r1 = (json.load(open(tmp_path / f"w2r{r}.json")) for r in range(2))
    assert r0["backend"] == "nccl" and {r0["device"], r1["device"]} == {0, 1}
    # member i starts on rank i mod 2; the worker then ships one member 0 -> 1 over NCCL before training, and the
    # Trainer's rebalancing may move more after each rung: wherever they live, all 8 are somewhere, once
    assert r0["local_members"] + r1["local_members"] == 8 and s["local_members"] == 8
    assert r0["migrations"] == r1["migrations"] >= 1 and s["migrations"] == 0
    assert not set(r0["steps"]) & set(r1["steps"])                                             # nobody trained twice
    assert r0["candidates"] == r1["candidates"]                                                # same decision everywhere
    assert r0["scores"] == r1["scores"] and len(r0["scores"]) == 8                             # every rank saw every score
    # placement changes nothing: same scores, same survivors, same step counts as one process holding all 8 members
    assert r0["candidates"] == s["candidates"]
    assert {**r0["steps"], **r1["steps"]} == s["steps"]
    for a, v in s["scores"].items():
        assert r0["scores"][a] == pytest.approx(v, abs=1e-6), a
    # save-then-resume in the middle of the second sweep, after members have moved: every rank rebuilds its population
    # (spare slots included), its experiments and the owner map from its own shard and ends where the uninterrupted job did
    again = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                            "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), worker, str(tmp_path), "8", "1"],
                           env=env, capture_output=True, text=True, timeout=600)
    assert again.returncode == 0, again.stdout[-2000:] + again.stderr[-2000:]
    q0, q1 = (json.load(open(tmp_path / f"w2r{r}_resumed.json")) for r in range(2))
    assert q0["candidates"] == q1["candidates"] == r0["candidates"]
    assert q0["steps"] == r0["steps"] and q1["steps"] == r1["steps"]            # same members in the same places
    for a, v in r0["scores"].items():
        assert q0["scores"][a] == pytest.approx(v, abs=1e-6) and q1["scores"][a] == pytest.approx(v, abs=1e-6), a
