"""Multi-GPU (needs >= 2 devices; skipped on the one-GPU box): the pattern plan over subject-partitioned stores with
rfgpu_shuffle (partition kernel + NCCL all-to-all) between the joins, identical to the single-store oracle."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_chain_join_with_nccl_shuffle(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29542", os.path.join(ROOT, "tests", "dist_worker.py"), "gpu", "200000"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for name in ("chain", "optional", "star"):
        assert f"world={world} {name}:" in r.stdout, r.stdout
