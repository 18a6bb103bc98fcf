"""Multi-GPU (needs >= 2 devices; skipped on the one-GPU box): the pattern plan over subject-partitioned stores with
rfgpu_shuffle between the joins, identical to the single-store oracle — through the fused partition + peer-memory scatter
kernel (default), the NCCL exchange (option off) and the overflow fallback (arena too small)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_world(world, exchange, n_triples, layout="subject"):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29542", os.path.join(ROOT, "tests", "dist_worker.py"), "gpu", str(n_triples),
           exchange, layout]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for name in ("chain", "optional", "star"):
        assert f"layout={layout} world={world} {name}:" in r.stdout, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_chain_join_with_p2p_shuffle(world):
    run_world(world, "p2p", 200000)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_chain_join_with_nccl_shuffle(world):
    run_world(world, "nccl", 200000)


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["hash", "range"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_chain_join_over_sharded_stores(world, layout):
    """one shard per permutation (rfgpu_store_set_sharded): quads placed by the hash of the component a permutation is joined
    on, or contiguous row slices of the globally sorted permutation (range partitioning by leading key)"""
    run_world(world, "p2p", 200000, layout)


@pytest.mark.gpu
def test_shuffle_arena_overflow_falls_back_on_every_rank():
    run_world(2, "tiny", 600000)
