"""imagenet-resnet50-multiworkers.py, torch flavour: one process per GPU ("8 localhost workers",
BASELINE.json configs[3]) -- the MultiWorkerMirroredStrategy shape (:20-25).

    strategy = tf.distribute.MultiWorkerMirroredStrategy(cluster_resolver=..., communication_options=NCCL)
becomes
    cross_ops = B200CrossDeviceOps(rank=task_id, nranks=num_workers, device=local_gpu)
    cross_ops.swap_into(strategy)        # MWMS takes no cross_device_ops argument

Each worker holds ONE replica.  Per step the 214 gradients go through
``batch_reduce(ReduceOp.SUM, ...)`` on the rank-context CrossDeviceOps (what CollectiveAllReduce
does under `model.fit`, :95) and a stock optimizer applies them; ``--fused`` uses the fused
reduce-scatter + Adam + all-gather step instead (one kernel, sharded optimizer state).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 examples/resnet50_multiworkers_torch.py
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import harness  # noqa: E402
from b200dp.cross_device_ops import B200CrossDeviceOps, CommunicationOptions, ReduceOp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=128, help="per replica (:70)")
    ap.add_argument("--fused", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))                 # the cluster resolver's task id (:16)
    n = int(os.environ.get("WORLD_SIZE", "1"))              # n_workers = SLURM_NTASKS (:29)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if n > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))   # rendezvous only
    print("Number of workers: ", n)
    torch.backends.cudnn.allow_tf32 = True
    torch.manual_seed(42)                                    # with strategy.scope(): identical initial variables
    model = harness.KerasResNet50().cuda().to(memory_format=torch.channels_last).train()
    params = model.keras_order_params()
    options = CommunicationOptions(implementation="NCCL")    # :20-21 (bytes_per_pack=0: one pack)
    cross_ops = B200CrossDeviceOps(rank=rank, nranks=n, device=local)
    g = torch.Generator(device="cuda").manual_seed(100 + rank)       # AutoShardPolicy.DATA (:66-69)
    x = (torch.rand(args.batch, 3, 224, 224, generator=g, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (args.batch,), generator=g, device="cuda")
    if args.fused:
        plan = b200dp.Plan(b200dp.Context.rank(rank, n, local), b200dp.Layout([p.numel() for p in params], nranks=n))
        if n > 1:
            obj = ["mw-%s" % os.environ.get("MASTER_PORT", "0")]
            dist.broadcast_object_list(obj, src=0)
            plan.exchange(obj[0])
            dist.barrier()
        stepper = harness.DataParallelStep(model, plan, optimizer="adam", lr=1e-3)
    else:
        opt = torch.optim.Adam(params, lr=1e-3, eps=1e-7)
    for step in range(args.steps):
        if args.fused:
            loss = stepper(x, y)
        else:
            for p in params:
                p.grad = None
            loss = F.cross_entropy(model(x), y) / n          # Keras scales the per-replica loss by 1/n
            loss.backward()
            # Keras under MWMS: batch_reduce(SUM, [(PerReplica grad_k, variable_k)] * 214); one local replica
            pairs = [([p.grad.contiguous()], None) for p in params]
            reduced = cross_ops.batch_reduce(ReduceOp.SUM, pairs, options)
            for p, r in zip(params, reduced):
                p.grad = r[0].view_as(p)
            opt.step()
        # a metric, the way Keras reduces it in replica context (newer TF: CrossDeviceOps._all_reduce)
        mean_loss = cross_ops._all_reduce(ReduceOp.SUM, loss.detach().reshape(1), 0, options)
        if rank == 0:
            print("step %d  global-batch loss %.4f" % (step + 1, float(mean_loss)))
    gathered = cross_ops.gather([torch.full((2,), float(rank), device="cuda")], axis=0)[0]
    assert gathered.tolist() == [float(r) for r in range(n) for _ in range(2)], gathered
    torch.cuda.synchronize()
    if n > 1:
        flat = torch.cat([p.detach().flatten() for p in params])
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        print("worker %d holds the same weights as worker 0: %s" % (rank, bool(torch.equal(flat, ref))))
        dist.barrier()
        dist.destroy_process_group()
    print("All done for worker", rank)


if __name__ == "__main__":
    main()
