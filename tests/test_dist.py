"""One process per rank over torch.distributed (tests/dist_worker.py).

not gpu : gloo, world_size 2 and 3 -- the N>1 host logic (halo plan, block split,
          read_petsc, create_A) over the multi-process transport, bit-exact against the
          reference's own code (oracle/_ref) run with the same number of thread-ranks.
gpu     : nccl, one rank per visible GPU (world_size 1 on a single-GPU box) -- update +
          mult + cg through the whole stack against the same oracle.
"""
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "dist_worker.py")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(mode, nproc, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER, mode]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    res = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=timeout)
    assert res.returncode == 0 and "DIST_OK" in res.stdout, res.stdout[-4000:]


@pytest.mark.parametrize("nproc", [2, 3, 5])
def test_host_logic_gloo(nproc):
    _launch("cpu", nproc)


@pytest.mark.gpu
def test_full_stack_nccl():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    _launch("gpu", max(1, min(torch.cuda.device_count(), 8)))
