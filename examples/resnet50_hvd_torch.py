"""imagenet-resnet50-hvd.py, torch flavour: one process per GPU (launch with torchrun or mpirun),
``import b200dp.hvd as hvd`` in place of ``import horovod.tensorflow.keras as hvd`` (:15).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 examples/resnet50_hvd_torch.py
    python examples/resnet50_hvd_torch.py                                   # size 1
"""
import argparse
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp import harness  # noqa: E402
from b200dp import hvd  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32)          # batch_size = 32 (:25)
    args = ap.parse_args()
    hvd.init()                                                 # :16
    torch.cuda.set_device(hvd.local_rank())                    # :36-41
    torch.backends.cudnn.allow_tf32 = True
    torch.manual_seed(hvd.rank())                              # ranks start different on purpose ...
    model = harness.KerasResNet50().cuda().to(memory_format=torch.channels_last).train()
    params = model.keras_order_params()
    # opt = hvd.DistributedOptimizer(Adam(lr)) (:99-101), with Adam fused into the reduction
    opt = hvd.FusedDistributedAdam(params, lr=1e-3)
    opt.broadcast_weights(root_rank=0)                         # ... BroadcastGlobalVariablesCallback(0) (:111)
    g = torch.Generator(device="cuda").manual_seed(100 + hvd.rank())    # dataset.shard(size, rank) (:78)
    x = (torch.rand(args.batch, 3, 224, 224, generator=g, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (args.batch,), generator=g, device="cuda")
    start = time.time()                                        # :119
    for step in range(args.steps):
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        avg = hvd.allreduce(loss.detach().reshape(1), op=hvd.Average)      # MetricAverageCallback (:113)
        if hvd.rank() == 0:                                    # verbose only on rank 0 (:117)
            print("step %d  mean loss over %d ranks %.4f" % (step + 1, hvd.size(), float(avg)))
    torch.cuda.synchronize()
    if hvd.rank() == 0:
        print("Total time: ", round(time.time() - start, 2), "(s)")       # :126
    opt.plan.check(0)
    print("All done for rank", hvd.rank())                     # :131
    if hvd.size() > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
