"""The example scripts (torch flavours of imagenet-resnet50-mirror.py and -hvd.py) run end to end."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def run(cmd, timeout=600):
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, (out.stdout[-1500:], out.stderr[-3000:])
    return out.stdout


def test_hvd_example_single_process():
    out = run([sys.executable, "examples/resnet50_hvd_torch.py", "--steps", "2", "--batch", "4"])
    assert "All done for rank 0" in out and "mean loss over 1 ranks" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_mirror_example_all_gpus():
    out = run([sys.executable, "examples/resnet50_mirror_torch.py", "--steps", "2", "--batch", "4", "--gpus", "2"])
    assert "replicas hold identical weights: True" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_hvd_example_two_ranks_torchrun():
    out = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
               "127.0.0.1", "--master-port", "29577", "examples/resnet50_hvd_torch.py", "--steps", "2", "--batch", "4"])
    assert "mean loss over 2 ranks" in out and "All done for rank 1" in out
