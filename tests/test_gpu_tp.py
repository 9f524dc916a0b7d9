"""Tensor-parallel shards (b200_tp_init + NCCL all-reduce) against the CPU oracle: needs >= 2 GPUs, otherwise skipped.
Runs tools/tp_check.py under torchrun, one rank per GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2])
def test_tensor_parallel_matches_oracle(world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(ROOT, "tools", "tp_check.py")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "TP CHECK PASS" in r.stdout


def test_tp_world_one_is_a_no_op():
    import nats_llm_studio_b200 as b200
    eng = b200.Engine(max_ctx=2048)
    assert len(eng.tp_unique_id()) == 128
