#!/usr/bin/env python
"""bench.py -- ResNet-50 data-parallel training throughput with the B200-native hot path.

Metric (BASELINE.json): ResNet-50 images/sec at 1/2/4/8 B200, with the per-step gradient
all-reduce + optimizer apply (the hot path) going through libb200dp.so; the same line also
carries the hot path's own time, its all-reduce bus bandwidth and the roofline of its
dominant kernel.

  python bench.py [--gpus N] [--steps K] [--warmup W]            one JSON line
  torchrun ... bench.py --gpus N ...                              one rank per GPU
  python bench.py --impl reference ...                            CPU restatement arm
  python bench.py --hotpath-only ...                              skip the conv graph (ncu runs)

A step = forward/backward of the Keras-shaped ResNet-50 on one synthetic batch (stand-in
torch/cuDNN graph: TensorFlow is absent on this image) + pack + fused
reduce-scatter/Adam/all-gather.  `value` has the batch resident in HBM; `e2e` copies every
step's batch from pinned host memory and reads the loss back, inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "resnet50_train_images_per_sec"
UNIT = "images/s"
N_ELEMS = 25_583_592          # Keras ResNet-50 trainable elements
NVLINK_PEAK_GBS = 770.0       # measured peer copy per direction (B200_PROFILING.md); 900 nominal


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200dp", choices=["b200dp", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="images per GPU (BASELINE.json configs[1])")
    ap.add_argument("--optimizer", default="adam", choices=["adam", "sgdm"])
    ap.add_argument("--wire", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--path", default="auto", choices=["auto", "peer", "multicast"],
                    help="N>1 collective path: peer loads with rank-order sum (bit-exact) or NVSwitch multicast")
    ap.add_argument("--overlap", action="store_true",
                    help="issue each bucket's pack+step from autograd hooks while backward still runs (SURVEY 8f-1)")
    ap.add_argument("--bucket-mb", type=float, default=0.0, help="bucket size in MB (0 = one bucket, the reference default)")
    ap.add_argument("--hotpath-only", action="store_true", help="time only pack+step on synthetic gradients")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed parity leg (tuning runs)")
    ap.add_argument("--parity-dist", default="D1", choices=["D1", "D2", "D3"])
    ap.add_argument("--ref-batch", type=int, default=32, help="CPU arm: images per step (configs[0])")
    return ap.parse_args()


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            try:
                self.proc.kill()
            except Exception:
                pass
        try:
            sm, reasons, mx = [], set(), None
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx = float(f[2])
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                     f[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
            if sm:
                sm.sort()
                out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))
        except Exception:
            pass
        return out


# ----------------------------------------------------------------------------- CPU arm
def cpu_training_step_baseline(n_replicas, batch, steps, warmup):
    """The reference's CPU path restated (imagenet-resnet50.py, and MirroredStrategy with CPU
    replicas): every step runs forward/backward of the Keras-shaped ResNet-50 for EACH of the
    n_replicas on the host cores, one replica after the other (torch CPU; TensorFlow is absent),
    then the oracle's replica-order AddN over the n gradient lists and TF-form Adam over all
    25.58M elements.  The batch is fixed: `batch` images per replica per step, a bounded sample
    of the 256 images per replica of the GPU arm's configuration (nothing shrinks itself)."""
    import numpy as np
    import torch
    import oracle
    from b200dp import harness
    cores = host_cores()
    torch.set_num_threads(cores)
    oracle.set_threads(cores)
    model = harness.KerasResNet50().train()
    params = model.keras_order_params()
    g = torch.Generator().manual_seed(7)
    xs = [torch.rand(batch, 3, 224, 224, generator=g) * 255.0 for _ in range(n_replicas)]
    ys = [torch.randint(0, 1000, (batch,), generator=g) for _ in range(n_replicas)]
    N = sum(p.numel() for p in params)
    w = np.concatenate([p.detach().numpy().ravel() for p in params]).astype(np.float32)
    m, v = np.zeros(N, np.float32), np.zeros(N, np.float32)
    gsum = np.empty(N, np.float32)
    times, hot = [], []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        grads = []
        for r in range(n_replicas):
            for p in params:
                p.grad = None
            loss = torch.nn.functional.cross_entropy(model(xs[r]), ys[r]) / n_replicas
            loss.backward()
            grads.append(np.concatenate([p.grad.numpy().ravel() for p in params]))
        t1 = time.perf_counter()
        oracle.addn(grads, out=gsum)
        oracle.adam_(w, m, v, gsum, lr=1e-3, step=it + 1)
        t2 = time.perf_counter()
        o = 0
        with torch.no_grad():
            for p in params:
                p.copy_(torch.from_numpy(w[o:o + p.numel()]).view(p.shape))
                o += p.numel()
        t3 = time.perf_counter()
        if it >= warmup:
            times.append(t3 - t0)
            hot.append(t2 - t1)
    dt = sum(times) / len(times)
    return {"images_per_s": n_replicas * batch / dt, "ms_per_step": dt * 1e3, "hot_path_ms": sum(hot) / len(hot) * 1e3,
            "cores": cores, "loss": float(loss.detach()), "batch": batch, "replicas": n_replicas}


def cpu_sample_text(steps, warmup, batch, n):
    return ("%d timed steps (%d warm-up), each step = %d replica(s) x %d images (a bounded sample of the 256 images per "
            "replica of the GPU arm's workload) fwd+bwd on torch-CPU, then oracle AddN over the %d replica gradient "
            "lists + TF-form Adam over all 25,583,592 elements; a port of the reference's CPU path "
            "(imagenet-resnet50.py / MirroredStrategy on CPU devices): TensorFlow itself is not installable here"
            % (steps, warmup, n, batch, n))


def clean_cpu_env():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the CPU arm is meant to use every
    host core, so it runs in a child process without the launcher's variables."""
    env = dict(os.environ)
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE",
              "GROUP_RANK", "ROLE_RANK", "TORCHELASTIC_RUN_ID", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    env["B2DP_CPU_CHILD"] = "1"
    env["CUDA_VISIBLE_DEVICES"] = ""
    return env


def cpu_baseline_subprocess(n, batch, steps, warmup):
    """cpu_training_step_baseline in a fresh interpreter (all host cores, no CUDA context, none of the
    launcher's thread limits); returns its result dict."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--gpus", str(n), "--steps", str(steps),
           "--warmup", str(warmup), "--ref-batch", str(batch)]
    out = subprocess.run(cmd, capture_output=True, text=True, env=clean_cpu_env(), cwd=ROOT, timeout=3000)
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    if out.returncode != 0 or not lines:
        raise RuntimeError("CPU baseline child failed: rc %d\n%s" % (out.returncode, out.stderr[-2000:]))
    return json.loads(lines[-1])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if os.environ.get("B2DP_CPU_CHILD") != "1" and (os.environ.get("OMP_NUM_THREADS") or "TORCHELASTIC_RUN_ID" in os.environ):
        line = cpu_baseline_subprocess(args.gpus, args.ref_batch, args.steps, args.warmup)   # same arm, clean environment
        print(json.dumps(line))
        return
    steps, warmup = max(1, args.steps), max(1, args.warmup)
    # the same `config` as the GPU arm of this N (what the CPU arm samples of it is in cpu_baseline.sample)
    if args.gpus == 1:
        args.path_used = "n/a (single GPU)"
    else:
        args.path_used = args.path if args.path != "auto" else ("multicast" if args.gpus > 4 else "peer")
    r = cpu_training_step_baseline(args.gpus, args.ref_batch, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["images_per_s"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": r["images_per_s"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": cpu_sample_text(steps, warmup, args.ref_batch, args.gpus),
                         "images_per_step": args.gpus * args.ref_batch, "hot_path_ms": r["hot_path_ms"],
                         "conv_graph": "torch-CPU stand-in"},
        "e2e": {"value": r["images_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, n):
    """The workload both arms are quoted on (BASELINE.json configs[1]); what the CPU arm samples of
    it is stated in its cpu_baseline.sample."""
    return {
        "workload": "imagenet-resnet50-mirror.py: Keras ResNet-50 (214 tensors, 25,583,592 params), synthetic "
                    "224x224x3, batch %d/GPU, Adam; gradient SUM + optimizer apply via b200dp" % args.batch,
        "global_batch": args.batch * n, "per_gpu_batch": args.batch,
        "optimizer": args.optimizer, "wire_dtype": args.wire, "parallelism": "dp%d" % n,
        "collective_path": getattr(args, "path_used", "n/a"),
        "buckets": getattr(args, "num_buckets", 1), "overlap_with_backward": bool(args.overlap),
        "conv_graph": "stand-in for the TF/cuDNN graph (TensorFlow absent): torch+cuDNN, TF32 convs as TF defaults, on the "
                      "GPU arm; torch-CPU on the reference arm",
        "l2": "inputs_exceed_l2 (activations and 409 MB of w/m/v/g per step >> 126 MB L2)",
    }


# ----------------------------------------------------------------------------- GPU arm
class NvlinkCounters:
    """Per-GPU NVLink data bytes (all links) from NVML's throughput counters: what actually crossed
    the wire around a timed loop, next to the algorithmic figure the roofline uses."""
    TX, RX, ALL_LINKS = 138, 139, 0xFFFFFFFF   # NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_{TX,RX}, KiB

    def __init__(self, gpu_index):
        self.h, self.why = None, None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.read()
        except Exception as e:   # counters are evidence, not the product: report why they are missing
            self.h, self.why = None, "%s: %s" % (type(e).__name__, e)

    def read(self):
        vals = self.nv.nvmlDeviceGetFieldValues(self.h, [(self.TX, self.ALL_LINKS), (self.RX, self.ALL_LINKS)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                raise RuntimeError("nvml field %d returned %d" % (v.fieldId, v.nvmlReturn))
            out.append(int(v.value.ullVal) * 1024)
        return out   # [tx_bytes, rx_bytes]


def parity_leg(args, plan, rank, local_rank, n):
    """Untimed, before anything is measured: 3 Adam steps with the lr schedule [1e-3,1e-3,1e-4] through
    pack_step on the collective path this line reports, on the seeded D1 gradients of SURVEY 8d, against
    the CPU oracle -- bit-exact on the peer path, inside the per-element interval bound on the multicast
    path -- plus NCCL's all_reduce of the same gradients inside the order-free bound."""
    import torch
    import torch.distributed as dist
    from oracle import parity
    try:
        res = parity.run(plan, [(0, rank, local_rank)], dist_name=args.parity_dist, nccl=(n > 1),
                         threads=max(1, host_cores() // n))
    except Exception as e:   # a crash in the checker is a failed check, reported as such
        import traceback
        res = {"checked": True, "ok": False, "error": "%s: %s" % (type(e).__name__, e), "trace": traceback.format_exc()[-800:]}
    if n > 1:
        allres = [None] * n
        dist.all_gather_object(allres, res)
    else:
        allres = [res]
    merged = dict(allres[0])
    merged["ok"] = all(r.get("ok") for r in allres)
    for key in ("max_abs_err", "max_err_over_bound", "nccl_max_err_over_bound", "w_max_bound", "w_median_bound"):
        vals = [r[key] for r in allres if key in r]
        if vals:
            merged[key] = max(vals)
    if "bit_exact" in merged:
        merged["bit_exact"] = all(r.get("bit_exact") for r in allres)
    if "nccl_within_bound" in merged:
        merged["nccl_within_bound"] = all(r.get("nccl_within_bound") for r in allres)
    if "w_frac_bit_equal" in merged:
        merged["w_frac_bit_equal"] = min(r["w_frac_bit_equal"] for r in allres if "w_frac_bit_equal" in r)
    errs = [r["error"] for r in allres if "error" in r]
    if errs:
        merged["error"] = errs[0]
    merged["ranks_checked"] = len(allres)
    # leave the plan as a fresh one: zero optimizer state (the stepper reloads the weights itself)
    import b200dp
    plan.tensor(0, b200dp.BUF_STATE0).zero_()
    plan.tensor(0, b200dp.BUF_STATE1).zero_()
    torch.cuda.synchronize()
    return merged


def run_b200dp(args):
    import torch
    import torch.distributed as dist
    import b200dp
    from b200dp import harness

    n = args.gpus
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != n:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N>1 with torch.distributed.run, one rank per GPU" % (n, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200; the library has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if n > 1:
        dist.init_process_group("nccl", device_id=device)
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True

    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.rank(rank, n, local_rank)
    bucket_bytes = int(args.bucket_mb * (1 << 20)) or (25 << 20 if args.overlap else 0)
    lay = b200dp.Layout(numels, nranks=n, wire=args.wire, bucket_bytes=bucket_bytes)
    plan = b200dp.Plan(ctx, lay)
    if n > 1:
        session = "bench-%s-%d" % (os.environ.get("MASTER_PORT", "0"), os.getppid())
        obj = [session]
        dist.broadcast_object_list(obj, src=0)
        plan.exchange(obj[0])
        dist.barrier()
        # auto: the multicast path only pays from 5 ranks up (r01 sweep: n=2 peer 176 vs mc 274 us,
        # n=4 251 vs 247, n=8 294 vs 241); below that the bit-exact peer path is also the faster one
        want_mc = args.path == "multicast" or (args.path == "auto" and n > 4)
        if want_mc:
            try:
                plan.enable_multicast(obj[0])
            except b200dp.B200dpError:
                if args.path == "multicast":
                    raise
        dist.barrier()

    args.path_used = plan.path if n > 1 else "n/a (single GPU)"
    args.num_buckets = lay.num_buckets
    parity_res = None
    if not args.no_parity:
        parity_res = parity_leg(args, plan, rank, local_rank, n)
        if not parity_res["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "value": None, "unit": UNIT, "n_gpus": n, "parity": parity_res,
                                  "error": "parity check against the oracle FAILED; nothing was timed"}))
            if n > 1:
                dist.barrier()
                dist.destroy_process_group()
            sys.exit(3)
    B = args.batch
    gen = torch.Generator(device=device).manual_seed(7 + rank)
    host_gen = torch.Generator().manual_seed(7 + rank)
    launches_per_step = 1 if (n == 1 and args.wire == "f32") else 2  # n=1: one fused kernel; n>1: pack + fused step

    if args.hotpath_only:
        flat_g = torch.randn(sum(numels), generator=gen, device=device) * 1e-3
        grads, o = [], 0
        for k in numels:
            grads.append(flat_g[o:o + k])
            o += k
        plan.tensor(0, b200dp.BUF_WEIGHTS).uniform_(-0.05, 0.05)
        t = [0]

        gptrs = b200dp.ptr_array(grads)

        def hot():
            t[0] += 1
            plan.pack_step(0, gptrs, args.optimizer, b200dp.adam_hyper(step=t[0]) if args.optimizer == "adam"
                           else b200dp.sgdm_hyper())

        def step_resident():
            hot()
            return None

        step_e2e = None
        stepper = None
    else:
        model = harness.KerasResNet50().to(device).to(memory_format=torch.channels_last).train()
        stepper = harness.DataParallelStep(model, plan, dev=0, optimizer=args.optimizer, lr=1e-3, overlap=args.overlap)
        x_dev = (torch.rand(B, 3, 224, 224, generator=gen, device=device) * 255.0).contiguous(
            memory_format=torch.channels_last)
        y_dev = torch.randint(0, 1000, (B,), generator=gen, device=device)
        # host batch as tf.data hands it over: float32 NHWC (the scripts cast to float32, -mirror.py:49)
        x_host = (torch.rand(B, 224, 224, 3, generator=host_gen) * 255.0).pin_memory()
        y_host = torch.randint(0, 1000, (B,), generator=host_gen).pin_memory()
        # double-buffered staging: the NEXT step's batch is copied on a side stream while this
        # step computes (what tf.data's prefetch-to-device does); every step still copies its own
        # batch inside the timed region
        x_stage = [torch.empty(B, 224, 224, 3, device=device) for _ in range(2)]
        y_stage = [torch.empty(B, dtype=torch.long, device=device) for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=device)
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        loss_host = torch.empty((), pin_memory=True)
        e2e_state = {"i": 0, "primed": False}

        def prefetch(slot):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[slot])
                x_stage[slot].copy_(x_host, non_blocking=True)
                y_stage[slot].copy_(y_host, non_blocking=True)
                ready[slot].record(copy_stream)

        def step_resident():
            return stepper(x_dev, y_dev)

        def step_e2e():
            cur = e2e_state["i"] & 1
            if not e2e_state["primed"]:
                consumed[0].record()
                consumed[1].record()
                prefetch(cur)
                e2e_state["primed"] = True
            torch.cuda.current_stream().wait_event(ready[cur])
            loss = stepper(x_stage[cur].permute(0, 3, 1, 2), y_stage[cur])   # NHWC memory viewed as channels_last NCHW
            consumed[cur].record()
            prefetch(cur ^ 1)                                  # next step's H2D overlaps this step's compute
            loss_host.copy_(loss.detach(), non_blocking=True)
            torch.cuda.current_stream().synchronize()          # the user reads the loss every step
            e2e_state["i"] += 1
            return loss_host

        hot = stepper.hot_path

    def barrier():
        if n > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, hot_events=False):
        for _ in range(warmup):
            fn()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        if n > 1:
            tms = torch.tensor([ms], device=device)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms / steps

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_step = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None

    # the hot path alone, and its dominant kernel alone, timed with events on the launch stream
    # (warm caches are not an issue: 409 MB of w/m/v/g per rank per step >> L2)
    def time_region(fn, reps):
        barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        torch.cuda._sleep(int(2e7))   # ~10 ms: lets the host queue every launch, so events time the GPU only
        e[0].record()
        for i in range(reps):
            fn()
            e[i + 1].record()
        barrier()
        ts = sorted(e[i].elapsed_time(e[i + 1]) for i in range(reps))
        t = torch.tensor([ts[len(ts) // 2], sum(ts) / len(ts)], device=device)
        if n > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    grads_for_pack = b200dp.ptr_array(grads if args.hotpath_only else [p.grad for p in stepper.params])
    tcount = [10 ** 6]

    def only_step():
        tcount[0] += 1
        h = b200dp.adam_hyper(step=tcount[0]) if args.optimizer == "adam" else b200dp.sgdm_hyper()
        plan.step(0, args.optimizer, h)

    def only_pack():
        plan.pack(0, grads_for_pack)

    def pack_and_step():          # the per-step call a binding makes (one kernel at n == 1)
        tcount[0] += 1
        h = b200dp.adam_hyper(step=tcount[0]) if args.optimizer == "adam" else b200dp.sgdm_hyper()
        plan.pack_step(0, grads_for_pack, args.optimizer, h)

    # the same call when the framework hands over gradient buffers at NEW addresses every step (what
    # torch's `p.grad = None` loop does): the library then re-sends its 214-entry pointer table and, at
    # n == 1, re-resolves the tile table (resolve_tiles_kernel) inside the call
    src_grads = grads if args.hotpath_only else [p.grad for p in stepper.params]
    alt_flat = torch.empty(sum(numels) + 64, device=device)
    alt, o = [], 16                         # a second copy of the gradients at different (still 64-byte aligned) addresses
    for gsrc in src_grads:
        alt.append(alt_flat[o:o + gsrc.numel()].copy_(gsrc.reshape(-1)))
        o += gsrc.numel()
    ptr_sets = [grads_for_pack, b200dp.ptr_array(alt)]

    def pack_and_step_new_ptrs():
        tcount[0] += 1
        h = b200dp.adam_hyper(step=tcount[0]) if args.optimizer == "adam" else b200dp.sgdm_hyper()
        plan.pack_step(0, ptr_sets[tcount[0] & 1], args.optimizer, h)

    reps = 30
    for _ in range(5):
        pack_and_step()
    hot_med_ms, hot_avg_ms = time_region(pack_and_step, reps)
    refresh_med_ms, _ = time_region(pack_and_step_new_ptrs, reps)
    pack_and_step()                          # back to the stable table
    nvl = NvlinkCounters(local_rank) if n > 1 else None
    nvl0 = nvl.read() if (nvl and nvl.h) else None
    step_med_ms, step_avg_ms = time_region(only_step, reps)
    nvl_traffic = None
    if nvl0 is not None:
        nvl1 = nvl.read()
        mine = torch.tensor([(nvl1[0] - nvl0[0]) / reps, (nvl1[1] - nvl0[1]) / reps], device=device, dtype=torch.float64)
        allv = [torch.zeros_like(mine) for _ in range(n)]
        dist.all_gather(allv, mine)
        nvl_traffic = {"source": "NVML NVLINK_THROUGHPUT_DATA_TX/RX, all links, around the %d-launch step loop" % reps,
                       "tx_bytes_per_launch_per_gpu": [float(v[0]) for v in allv],
                       "rx_bytes_per_launch_per_gpu": [float(v[1]) for v in allv]}
    elif nvl is not None:
        nvl_traffic = {"source": "NVML NVLink counters unavailable: %s" % nvl.why}
    sep_step_ms = step_avg_ms
    pack_med_ms, pack_avg_ms = time_region(only_pack, reps)
    # the same per-step call once more after the board has idled for a second: directly behind
    # ~800 W of convolution work the memory system runs a few percent slower (tools/investigate_n1.py)
    time.sleep(1.0)
    idle_med_ms, _ = time_region(pack_and_step, reps)
    plan.check(0)

    ar_med_ms = None
    if n > 1:
        only_pack()
        ar_med_ms, _ = time_region(lambda: plan.allreduce(0, postscale=1.0 / n), reps)
        plan.check(0)

    other = {}
    if n > 1 and plan.path == "multicast":   # the deterministic peer path on the same buffers, for the record
        plan.set_path("peer")
        for _ in range(3):
            pack_and_step()
        other["peer_path_step_us"] = time_region(only_step, reps)[1] * 1e3
        only_pack()
        other["peer_path_allreduce_only_us"] = time_region(lambda: plan.allreduce(0, postscale=1.0 / n), reps)[0] * 1e3
        plan.check(0)
        plan.set_path("multicast")

    # NCCL on the same box, same 102.33 MB fp32 message: the library call the reference's
    # strategies bottom out in, timed the same way (the bar to beat)
    nccl_med_ms = None
    if n > 1:
        buf = torch.randn(N_ELEMS, device=device) * 1e-3
        for _ in range(5):
            dist.all_reduce(buf)
        nccl_med_ms, _ = time_region(lambda: dist.all_reduce(buf), reps)
        del buf

    # b2dp_adam_apply_noalloc (the entry point for callers that cannot allocate) on the same 25.58M elements
    noalloc_ms = None
    if n == 1 and args.optimizer == "adam":
        na = [torch.zeros(N_ELEMS, device=device) for _ in range(3)]
        na_g = torch.randn(N_ELEMS, device=device) * 1e-3
        na_stats = torch.zeros(2, dtype=torch.float64, device=device)
        na_t = [0]

        def noalloc_call():
            na_t[0] += 1
            b200dp.adam_apply_noalloc(na[0], na[1], na[2], na_g, b200dp.adam_hyper(step=na_t[0]), na_stats)

        for _ in range(3):
            noalloc_call()
        noalloc_ms = time_region(noalloc_call, reps)[0]
        del na, na_g

    e2e = None
    if step_e2e is not None:
        ms_e2e = timed(step_e2e, args.steps, args.warmup)
        e2e = {"value": n * B / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": n * (x_host.numel() * 4 + y_host.numel() * 8), "d2h_bytes_per_step": n * 4}

    if rank != 0:
        if n > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    peaks, peak_kind = measured_peaks()
    elems_per_rank = lay.shard_elems          # padded slice; algorithmic bytes use the unpadded N/n
    S = 4 * N_ELEMS
    if n == 1:
        bytes_per_elem = 28 if args.optimizer == "adam" else 20
        alg_bytes = bytes_per_elem * N_ELEMS
        fused = args.wire == "f32"
        if fused:   # the dominant (only) kernel of the step is the in-place fused one
            step_avg_ms = hot_avg_ms
        achieved = alg_bytes / (step_avg_ms * 1e-3) / 1e9
        roofline = {"kernel": ("fused_tiles_kernel<%s> (optimizer apply reading the 214 gradients in place, 256-bit)" % args.optimizer) if fused
                    else "step_kernel<N=1,bf16,%s> (fused optimizer apply)" % args.optimizer,
                    "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm_gbs"], "peak_source": peak_kind + " (MEASURED_PEAKS.json copy bandwidth)",
                    "algorithmic_bytes_per_launch": alg_bytes, "us_per_launch": step_avg_ms * 1e3,
                    "traffic": load_traffic("fused_n1" if fused else "step_n1")}
    else:
        wire_b = 2 if args.wire == "bf16" else 4
        if plan.path == "multicast" and args.wire == "bf16":
            # busier direction = inbound: S replicated fp32 stores + S/(2n) reduced bf16; outbound S/2 + S/n
            nvl_bytes = S * (1.0 + 0.5 / n)
        elif plan.path == "multicast":   # inbound: S/n reduced + S replicated stores (the switch also writes the
            nvl_bytes = S * (1.0 + 1.0 / n)   # sender's own copy); outbound the mirror image: S fetched + S/n stored
        else:                          # per GPU per direction: RS reads in + AG writes in
            nvl_bytes = (n - 1) / n * (wire_b * N_ELEMS) + (n - 1) / n * S
        achieved = nvl_bytes / (step_avg_ms * 1e-3) / 1e9
        roofline = {"kernel": ("%s<%s> (multimem.ld_reduce + apply + multimem.st through NVSwitch, N=%d)"
                               % ("step_mc_bf16_kernel" if args.wire == "bf16" else "step_mc_kernel", args.optimizer, n)) if plan.path == "multicast" else
                              ("step256_kernel<N=%d,%s>" % (n, args.optimizer) if (args.wire == "f32" and n in (2, 4, 8))
                               else "step_kernel<N=%d,%s,%s>" % (n, args.wire, args.optimizer))
                              + " (reduce-scatter + apply + all-gather over NVLink peer memory)",
                    "bound": "nvlink", "achieved": achieved, "peak": NVLINK_PEAK_GBS, "unit": "GB/s",
                    "frac": achieved / NVLINK_PEAK_GBS,
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                    "algorithmic_bytes_per_launch": nvl_bytes, "us_per_launch": step_avg_ms * 1e3,
                    "traffic": (max(nvl_traffic["rx_bytes_per_launch_per_gpu"] + nvl_traffic["tx_bytes_per_launch_per_gpu"])
                                if nvl_traffic and "rx_bytes_per_launch_per_gpu" in nvl_traffic else None),
                    "traffic_note": "NVLink data bytes per launch on the busiest GPU and direction (NVML counters); "
                                    "algorithmic_bytes_per_launch is the per-direction payload",
                    "nvlink_counters": nvl_traffic}
    if n > 1 and plan.path == "peer":
        # what the same traffic pattern reaches with no arithmetic and no barriers when every GPU issues it at
        # once (tools/nvlink_ceiling.cu, committed result): the ceiling THIS kernel can be held to
        try:
            with open(os.path.join(ROOT, "profiles", "nvlink_ceiling.json")) as f:
                ceil = json.load(f).get(str(n))
            if ceil and args.wire == "f32":
                roofline["measured_pattern_ceiling_gbs"] = ceil["fused_pull_push_gbs_per_direction"]
                roofline["frac_of_measured_pattern_ceiling"] = achieved / ceil["fused_pull_push_gbs_per_direction"]
        except Exception:
            pass
    busbw = (S / (step_avg_ms * 1e-3)) * 2 * (n - 1) / n / 1e9 if n > 1 else 0.0
    hot_path = {"us_per_step_call_median": hot_med_ms * 1e3, "us_per_step_call_after_1s_idle": idle_med_ms * 1e3,
                "us_per_step_call_new_grad_addresses": refresh_med_ms * 1e3, "us_step_kernel_avg": step_avg_ms * 1e3,
                "us_separate_apply_kernel_avg": sep_step_ms * 1e3, "us_separate_pack_kernel_avg": pack_avg_ms * 1e3, "share_of_train_step": hot_avg_ms / ms_step,
                "allreduce_busbw_gbs": busbw, "busbw_vs_900": busbw / 900.0,
                "pack_hbm_gbs": (4 + (2 if args.wire == "bf16" else 4)) * N_ELEMS / (pack_avg_ms * 1e-3) / 1e9,
                "elems_per_rank": elems_per_rank}
    hot_path.update(other)
    if noalloc_ms is not None:
        hot_path["us_adam_apply_noalloc"] = noalloc_ms * 1e3
        hot_path["adam_apply_noalloc_hbm_gbs"] = 28 * N_ELEMS / (noalloc_ms * 1e-3) / 1e9
    if ar_med_ms:
        hot_path["allreduce_only_us"] = ar_med_ms * 1e3
        hot_path["allreduce_only_busbw_gbs"] = (S / (ar_med_ms * 1e-3)) * 2 * (n - 1) / n / 1e9
    if nccl_med_ms:
        hot_path["nccl_allreduce_us"] = nccl_med_ms * 1e3
        hot_path["nccl_busbw_gbs"] = (S / (nccl_med_ms * 1e-3)) * 2 * (n - 1) / n / 1e9

    line = {
        "metric": METRIC if not args.hotpath_only else "hot_path_steps_per_sec_as_images",
        "value": n * B / (ms_step * 1e-3), "unit": UNIT, "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, n),
        "clocks": clocks, "gpu_launches": launches_per_step * args.steps * lay.num_buckets,
        "gpu_launches_note": "libb200dp.so kernels per step: fused_tiles_kernel at n=1, pack_kernel + step kernel at n>1; "
                             "conv graph launches are torch/cuDNN's",
        "roofline": roofline, "hot_path": hot_path,
    }
    if e2e:
        line["e2e"] = e2e
    if parity_res is not None:
        line["parity"] = parity_res
    if n > 1:
        # every other rank leaves now: a rank parked in an NCCL barrier busy-waits on a host core, and one
        # busy core stalls every OpenMP region of the CPU baseline below (measured: 13x slower)
        dist.barrier()
        dist.destroy_process_group()
    if not args.no_cpu_baseline and not args.hotpath_only:
        cpu_steps = 2 if n == 1 else 1
        line["cpu_baseline"] = cpu_baseline_subprocess(n, args.ref_batch, cpu_steps, 1)["cpu_baseline"]
        # the one ratio this library owns: the reference's CPU AddN + Adam over the same n gradient lists
        # against the GPU per-step call (pack + fused reduce-scatter/Adam/all-gather)
        hot_path["hot_path_vs_cpu"] = line["cpu_baseline"]["hot_path_ms"] * 1e3 / hot_path["us_per_step_call_median"]
    print(json.dumps(line))


def load_traffic(key):
    """dram bytes (read+write) per launch from the committed ncu --set full capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(key)
    except Exception:
        return None


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200dp(a)
