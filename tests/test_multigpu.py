"""REAL multi-GPU runs of the multi-rank paths (NCCL + symmetric memory over NVLink): tests/multigpu_worker.py under
torch.distributed.run, one process per GPU.  Marker `multigpu` (and `gpu`): skipped on a box with fewer than 2 GPUs --
the one-GPU suite covers the same paths with emulated ranks (test_gpu_dist.py, test_gpu_shard_train.py), the CPU suite
covers the host logic over gloo (test_dist_cpu.py).  Logs of the 2- and 8-GPU runs: profiles/r2_multigpu_*.log."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # pragma: no cover
        return 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_real_ranks(world):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    port = 29600 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(REPO, "tests", "multigpu_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=REPO)
    sys.stdout.write(res.stdout)
    sys.stderr.write(res.stderr[-4000:])
    out_dir = os.path.join(REPO, "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "r2_multigpu_%d.log" % world), "w") as f:
            f.write(res.stdout + "\n---- stderr (tail) ----\n" + res.stderr[-8000:])
    assert res.returncode == 0 and "MULTIGPU CHECK OK" in res.stdout, res.stderr[-3000:]
