"""Multi-GPU: the target split over 2 (and 4) pipeline stages, one process per GPU over NCCL, must generate exactly what
one GPU generates, with identical KV bookkeeping on every stage.  Needs >= 2 GPUs on the box (gpurun --gpus 2)."""
import os
import subprocess
import sys

import pytest

from tests.test_pipeline_cpu import free_port

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


# head share 0: the head GPU keeps no target layer (it embeds the tokens and runs the draft), the reference's topology
@pytest.mark.parametrize("world,head_share", [(2, 0.25), (4, 0.25), (2, 0.0)])
def test_layer_split_target_matches_single_gpu(world, head_share):
    if n_gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    env = dict(os.environ, PIB_PIPE_TIMEOUT="120", PIB_TEST_HEAD_SHARE=str(head_share))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "tests", "pipeline_gpu_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-4000:]
    assert r.stdout.count("PIPELINE_GPU_OK") == world
