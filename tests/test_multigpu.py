"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): runs tests/mgpu_check.py under torchrun."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4])
def test_strip_partition_parity_on_gpus(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29600 + world), os.path.join(ROOT, "tests", "mgpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("MGPU_OK") == world, r.stdout[-3000:]


@pytest.mark.parametrize("world", [2, 4])
def test_peer_memory_path_equals_nccl_path(world, tmp_path):
    """The one-shot peer-memory all-reduce (rank-order sum) and mailbox halo exchange against NCCL: same iteration
    counts; the solution bitwise for 2 ranks (a + b has one rounding), to round-off beyond (NCCL's order is its own)."""
    import numpy as np
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    out = {}
    for p2p in ("1", "0"):
        env = dict(os.environ, PYN_P2P=p2p, MGPU_DUMP=str(tmp_path / f"p2p{p2p}"))
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
               "127.0.0.1", "--master-port", str(29620 + world + int(p2p)), os.path.join(ROOT, "tests", "mgpu_check.py")]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        assert r.stdout.count(f"peer_memory={p2p == '1'}") == world, r.stdout[-2000:]
        out[p2p] = [np.load(str(tmp_path / f"p2p{p2p}.{k}.npy")) for k in range(world)]
    for a, b in zip(out["1"], out["0"]):
        assert np.array_equal(a[:2], b[:2])
        if world == 2:
            assert np.array_equal(a, b)
        else:
            assert np.abs(a[2:] - b[2:]).max() <= 1e-12 * np.abs(b[2:]).max()
