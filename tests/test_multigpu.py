"""Multi-rank parity on real GPUs (needs >= 2 CUDA devices; skipped on a one-GPU box): tools/dp_check.py under
torchrun -- replicas bit-identical through the peer-memory exchange and through the NCCL formulation, the two
agreeing with each other, a prune + densify keeping the ranks in lockstep.  (The CPU counterpart for the torch path
is tests/test_mapper_gloo.py; the kernels' arithmetic is pinned on one device in tests/test_mapping.py.)"""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_ranks_stay_bit_identical_and_exchanges_agree():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29600 + os.getpid() % 300
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "dp_check.py")], capture_output=True, text=True, timeout=280)
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and line, (r.stdout[-2000:], r.stderr[-3000:])
    rep = json.loads(line[-1])
    assert rep["ok"] and rep["p2p_replicas_identical"] and rep["nccl_replicas_identical"], rep
    assert rep["p2p_used"] == "p2p", rep   # the peer-memory path really ran (no silent fallback)
