#!/bin/bash
# Developer sweep on N GPUs (default 8): hot path only, both collective paths and the tuning knobs.
N=${1:-8}
run() {
  echo "== $*"
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29533 bench.py --gpus $N --hotpath-only --steps 20 --warmup 5 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); h=d['hot_path']
print({k: round(v,1) for k,v in h.items() if k in ('us_step_kernel_avg','allreduce_only_us','peer_path_step_us','peer_path_allreduce_only_us','nccl_allreduce_us','allreduce_busbw_gbs')}, d['config']['collective_path'])"
}
run B2DP_MC_UNROLL=4
run B2DP_MC_UNROLL=2
run B2DP_MC_UNROLL=1
run B2DP_MC_UNROLL=4 B2DP_STEP_GRID_CAP=148
run B2DP_MC_UNROLL=2 B2DP_STEP_GRID_CAP=148
run B2DP_MC_UNROLL=4 B2DP_STEP_GRID_CAP=74
