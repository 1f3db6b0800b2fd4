"""The example scripts (torch flavours of imagenet-resnet50-mirror.py, -hvd.py, -multiworkers.py and -ps.py:
all five BASELINE.json configurations have a runnable shape) run end to end."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return str(port)


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
               "127.0.0.1", "--master-port", free_port(), "examples/resnet50_hvd_torch.py", "--steps", "2", "--batch", "4"])
    assert "mean loss over 2 ranks" in out and "All done for rank 1" in out


def test_multiworkers_example_single_worker():
    out = run([sys.executable, "examples/resnet50_multiworkers_torch.py", "--steps", "2", "--batch", "4"])
    assert "All done for worker 0" in out


def test_ps_example_single_gpu_saves_a_partitioned_checkpoint(tmp_path):
    path = str(tmp_path / "ps.npz")
    out = run([sys.executable, "examples/resnet50_ps_torch.py", "--steps", "2", "--batch", "4", "--gpus", "1", "--save", path])
    assert "workers hold identical weights: True" in out and "Saving model to" in out
    import numpy as np
    names = set(np.load(path).keys())
    assert "conv1_conv/kernel" in names and "optimizer/iterations" in names


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("fused", [False, True])
def test_multiworkers_example_two_workers(fused):
    out = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
               "127.0.0.1", "--master-port", free_port(), "examples/resnet50_multiworkers_torch.py", "--steps", "2",
               "--batch", "4"] + (["--fused"] if fused else []))
    assert "worker 1 holds the same weights as worker 0: True" in out and "All done for worker 1" in out


@pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")
def test_ps_example_two_gpus(tmp_path):
    path = str(tmp_path / "ps2.npz")
    out = run([sys.executable, "examples/resnet50_ps_torch.py", "--steps", "2", "--batch", "4", "--gpus", "2", "--save", path])
    assert "workers hold identical weights: True" in out
    import numpy as np
    names = set(np.load(path).keys())
    assert "dense/kernel/part_0" in names and "dense/kernel/part_1" in names and "optimizer/dense/kernel/m/part_1" in names
