"""-m gpu, needs >= 2 GPUs (skipped otherwise): ShardedFusedAdam (ghr_adam_step_sharded: reduce-scatter + Adam +
all-gather in one kernel over NVLink peer memory) equals NCCL all-reduce + FusedAdam, and keeps the replicas bit-identical."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_adam_matches_allreduce_plus_adam(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29530 + world),
                        os.path.join(ROOT, "tests", "dist", "sharded_adam_worker.py")],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and f"SHARDED_ADAM_OK {world}" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
