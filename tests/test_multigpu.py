"""Two GPUs, one process each (torchrun, NCCL): `gfalign_b200.multigpu` prints the reference's rows /
writes the reference's RC:i tags with the alignment table sharded over the ranks."""
import os
import subprocess
import sys

import pytest

from test_gpu_parity import _cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _torchrun(n, args, cwd, port, **extra_env):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), **extra_env)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), "-m", "gfalign_b200.multigpu"] + args
    return subprocess.run(cmd, cwd=cwd, env=env, capture_output=True, text=True, timeout=600)


def test_two_gpu_search_and_evalgfa_match_reference(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    port = 29610
    seen = 0
    for c in _cases():
        if c["name"] not in ("random3_search", "tangle1_search_all", "tangle2_search", "tangle2_evalgfa",
                             "hairpin_evalgfa"):
            continue
        d = os.path.join(str(tmp_path), c["name"])
        os.makedirs(d)
        for fn, txt in c["files"].items():
            with open(os.path.join(d, fn), "w") as f:
                f.write(txt)
        modes = (["--shard", "alignments"], ["--shard", "batch"]) if "_search" in c["name"] else ([],)
        for extra in modes:      # alignment shards and candidate-batch shards must print the same rows
            p = _torchrun(2, list(c["args"]) + extra, d, port)
            port += 1
            assert p.returncode == 0, p.stderr[-2000:]
            assert p.stdout == c["stdout"], (c["name"], extra)
        if "-o" in c["args"]:
            out = open(os.path.join(d, c["args"][c["args"].index("-o") + 1])).read()
            assert out == c["out_text"], c["name"]
        seen += 1
    assert seen == 5


def test_two_gpu_search_when_one_rank_has_no_incremental_path(tmp_path):
    """One rank without a pair dictionary (here forced: GFE_FORCE_LEGACY on rank 1 only) can only answer
    full scans; its peers must not keep sending incremental records under the same epochs.
    gfe_search_sharded votes once over the exchange buffer and all ranks fall back together: the rows stay
    the reference's."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    port = 29630
    seen = 0
    for c in _cases():
        if c["name"] not in ("tangle1_search_all", "tangle2_search"):
            continue
        d = os.path.join(str(tmp_path), c["name"])
        os.makedirs(d)
        for fn, txt in c["files"].items():
            with open(os.path.join(d, fn), "w") as f:
                f.write(txt)
        for legacy_rank in ("1", "0"):
            p = _torchrun(2, list(c["args"]) + ["--shard", "alignments"], d, port, GFE_TEST_LEGACY_RANK=legacy_rank,
                          GFE_PEER_TIMEOUT_MS="5000")
            port += 1
            assert p.returncode == 0, p.stderr[-2000:]
            assert p.stdout == c["stdout"], (c["name"], legacy_rank)
        seen += 1
    assert seen == 2


def test_peer_exchange_matches_nccl():
    """gfe_peer_sum_u32 (NVLink peer-memory sum of the count records) == NCCL all-reduce."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    n = min(torch.cuda.device_count(), 8)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29650",
           os.path.join(ROOT, "tests", "peer_exchange_worker.py")]
    p = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    assert "mismatching epochs 0" in p.stderr
