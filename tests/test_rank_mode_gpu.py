"""One process per GPU (torchrun + NCCL): the path bench.py and the driver's 1/2/4/8-GPU scaling run
take.  Spawns tools/rank_mode_check.py under torch.distributed.run with one rank per visible GPU and
requires every rank to report bit-identical results against the oracle.  A 1-GPU box still runs the
world-size-1 leg (rank-mode handles without a communicator)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tools", "rank_mode_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    for line in r.stdout.splitlines():   # kept in the log (pytest -s): which kernels ran on every rank
        if "RANK_MODE_OK" in line or "pair-symmetric cases" in line:
            print(line)
    for rank in range(world):
        assert f"RANK_MODE_OK rank {rank}/{world}" in r.stdout, r.stdout[-3000:]


def test_rank_mode_world_1(hgb):
    _run(1)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_rank_mode_nccl_vs_oracle(hgb, world):
    have = hgb.lib().hgb_device_count()
    if have < world:
        pytest.skip(f"needs {world} GPUs, box has {have}")
    _run(world)
