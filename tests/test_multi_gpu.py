"""One scene rendered by 2 (and, where present, 4) GPUs through bear_cuda_comm_connect equals the
one-GPU render and the oracle.  Needs >= 2 visible GPUs (skipped otherwise); the work is done by
tests/mp_shared_scene.py, one process per GPU under torch.distributed.run."""
import socket
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.timeout(900)
def test_shared_scene_matches_one_gpu_and_oracle(world):
    if _gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "mp_shared_scene.py")],
                       capture_output=True, text=True, timeout=850)
    sys.stdout.write(p.stdout[-3000:])
    sys.stderr.write(p.stderr[-3000:])
    assert p.returncode == 0 and "SHARED_SCENE_OK world=%d" % world in p.stdout
