"""Sharded parity on real GPUs: tools/multigpu_check.py (random circuits biased onto the shard-selecting qubits, sharded
density matrices with noise and measurement, the multi-register API fuzz, probability sweeps) under torchrun, once per
execution mode of the sharded engine.  Needs at least two GPUs; on a one-GPU box every case is skipped.  The host-side
logic of these modes is covered on CPU in tests/test_shard_plan.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MODES = {
    "default": {},
    "exchange_pipeline_forced_on_small_shards": {"QUEST_B200_OVERLAP_MIN_BITS": "8"},
    "exchanges_back_to_back": {"QUEST_B200_OVERLAP": "0"},
}


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("mode", sorted(MODES))
def test_sharded_parity(mode):
    n = _gpu_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    env = dict(os.environ, **MODES[mode])
    port = 29600 + sorted(MODES).index(mode)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "multigpu_check.py")]
    out = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)
    tail = "\n".join(out.stdout.splitlines()[-14:])
    assert out.returncode == 0 and "MULTIGPU PARITY PASS" in out.stdout, tail + "\n" + out.stderr[-2000:]
