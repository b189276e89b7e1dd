"""The population-sharded path on MORE THAN ONE GPU: scripts/check_popshard.py under torchrun shards one population over the
ranks through both transports (NCCL all-gather + merge kernels; key exchange fused into the iteration kernel over NVLink
peer memory) and compares every rank with the unsharded handle bit for bit.  Skipped on a one-GPU box."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def _n_gpus() -> int:
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("case", ["task", "sdf"])
def test_population_sharded_over_ranks_matches_unsharded(case):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = 2 if n < 4 else 4
    env = dict(os.environ, POP_PER_RANK="2048", POP_CASE=case, POP_SOAK="2000", MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", "29577", str(ROOT / "scripts" / "check_popshard.py")], env=env, capture_output=True, text=True, timeout=900)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    assert f"all-gather + merge: OK on {world} ranks" in out, out[-3000:]
    assert f"fused exchange: OK on {world} ranks" in out, out[-3000:]
    assert f"soak: OK on {world} ranks" in out, out[-3000:]
