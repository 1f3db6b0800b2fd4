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
    opt = hvd.FusedDistributedAdam(params, lr=0.1 * hvd.size())         # Adam(learning_rate=0.1*hvd.size()) (:99)
    steps_per_epoch = max(1, args.steps // 2)
    callbacks = [                                              # :108-115
        hvd.callbacks.BroadcastGlobalVariablesCallback(0),    # ... so that every rank starts from rank 0's weights
        hvd.callbacks.MetricAverageCallback(),
        hvd.callbacks.LearningRateWarmupCallback(0.1 * hvd.size(), warmup_epochs=3, steps_per_epoch=steps_per_epoch, verbose=1),
    ]
    for cb in callbacks:
        cb.set_optimizer(opt)
        cb.on_train_begin()
    g = torch.Generator(device="cuda").manual_seed(100 + hvd.rank())    # dataset.shard(size, rank) (:78)
    x = (torch.rand(args.batch, 3, 224, 224, generator=g, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (args.batch,), generator=g, device="cuda")
    start = time.time()                                        # :119
    for step in range(args.steps):                             # what model.fit drives (:120)
        epoch, batch = divmod(step, steps_per_epoch)
        for cb in callbacks:
            if batch == 0:
                cb.on_epoch_begin(epoch)
            cb.on_batch_begin(batch)
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        logs = {"loss": float(loss)}
        for cb in callbacks:
            cb.on_batch_end(batch, logs)
            if batch == steps_per_epoch - 1:
                cb.on_epoch_end(epoch, logs)                   # MetricAverageCallback: loss averaged over ranks (:113)
        if hvd.rank() == 0:                                    # verbose only on rank 0 (:117)
            print("step %d  lr %.4f  mean loss over %d ranks %.4f" % (step + 1, opt.lr, hvd.size(), logs["loss"]))
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
