"""Which NVLink byte counters does this box expose?  (developer probe; bench.py uses what works)"""
import subprocess

import pynvml as nv

nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(0)
for name, fid in (("DATA_TX", 138), ("DATA_RX", 139), ("RAW_TX", 140), ("RAW_RX", 141)):
    for scope in (0xFFFFFFFF, 0, 1):
        try:
            v = nv.nvmlDeviceGetFieldValues(h, [(fid, scope)])[0]
            print(name, "scope", hex(scope), "ret", v.nvmlReturn, "val", v.value.ullVal)
        except Exception as e:
            print(name, "scope", hex(scope), "EXC", e)
for fn in ("nvmlDeviceGetNvLinkUtilizationCounter",):
    try:
        print(fn, getattr(nv, fn)(h, 0, 0))
    except Exception as e:
        print(fn, "EXC", e)
for cmd in (["nvidia-smi", "nvlink", "-gt", "d", "-i", "0"], ["nvidia-smi", "nvlink", "-gt", "r", "-i", "0"],
            ["nvidia-smi", "nvlink", "-s", "-i", "0"]):
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=30)
        print("$", " ".join(cmd), "rc", out.returncode)
        print("\n".join((out.stdout + out.stderr).splitlines()[:12]))
    except Exception as e:
        print(cmd, "EXC", e)
try:
    import torch
    x = torch.zeros(1 << 26, device="cuda:0")
    y = torch.zeros(1 << 26, device="cuda:1")
    before = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", "0"], capture_output=True, text=True).stdout
    for _ in range(10):
        y.copy_(x)
    torch.cuda.synchronize()
    after = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", "0"], capture_output=True, text=True).stdout
    print("after 10 x 256 MiB peer copies 0->1:")
    print("\n".join(a + "   <-   " + b for a, b in list(zip(after.splitlines(), before.splitlines()))[:8]))
except Exception as e:
    print("copy probe EXC", e)
