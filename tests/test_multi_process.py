"""The one-process-per-GPU shape (cfg.rank / nranks / nccl_id, CUDA-IPC peer memory between the processes) -- what
bench.py and the driver's scaling run use, and the shape of nbody-par under MPI (nbody-par.c:775-779,913-917) --
checked bit for bit against a one-GPU run ON EVERY RANK.  Launched like the driver launches bench.py:
python -m torch.distributed.run --nproc-per-node P.  Needs >= 2 GPUs (skipped otherwise); bench.py --gpus N puts
the same check into the scaling run's JSON line (`parity`)."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _launch(world, *args):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd + [str(a) for a in args], capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-1500:] + r.stderr[-3000:]
    return json.loads(lines[-1])


def _worlds(nb):
    k = nb.device_count()
    if k < 2:
        pytest.skip("needs >= 2 GPUs")
    return [p for p in (2, 4, 8) if p <= k]


@pytest.mark.parametrize("p2p", [1, -1])
def test_symmetric_fast_sharded_processes_same_bits_as_one_gpu(nb, p2p):
    """Cell order + far field + symmetric kernel over P processes: peer-memory exchange (captured in a CUDA graph)
    and the NCCL exchange; a re-sort falls inside the run."""
    for world in _worlds(nb):
        out = _launch(world, "--bodies", 20000, "--mode", 1, "--p2p", p2p, "--k1", 5, "--k2", 4, "--sort-interval", 3)
        assert out["bits_equal_to_1gpu_on_every_rank"] is True, out
        assert ("peer-memory" in out["exchange"]) == (p2p == 1), out
        assert out["sha256_sharded"] == out["sha256_1gpu"]


def test_wide_kernel_sharded_processes_same_bits_as_one_gpu(nb):
    """N >= 49152: the 64 x 8 shape with short tail chunks, the configuration the headline number is made of."""
    for world in _worlds(nb):
        out = _launch(world, "--bodies", 131072 + 300, "--mode", 1, "--k1", 2, "--k2", 2)
        assert out["bits_equal_to_1gpu_on_every_rank"] is True, out
        assert out["kernel"] == "forces_sym_kernel<64,8>"


def test_exact_sharded_processes_same_bits_as_one_gpu(nb):
    """Exact mode, directed kernel, ncclAllGather of the new positions each step (any P)."""
    for world in _worlds(nb):
        out = _launch(world, "--bodies", 1001, "--mode", 0, "--k1", 6, "--k2", 3)
        assert out["bits_equal_to_1gpu_on_every_rank"] is True, out
        assert out["exchange"] == "ncclAllGather"
