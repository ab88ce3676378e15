"""CPU-only, world_size 2 and 3 over gloo: the host side of the multi-GPU pipeline (control ring, ggml_mpi_* shim, layer
partition, KV bookkeeping replicated to the other stages).  The ranks are tests/pipeline_cpu_worker.py."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_pipeline_host_logic_over_gloo(world):
    env = dict(os.environ, PIB_PIPE_TIMEOUT="60", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "tests", "pipeline_cpu_worker.py")]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("PIPELINE_CPU_OK") == world
