#!/bin/bash
# Round-2 measurement session on an 8-GPU box (gpurun --gpus 8): NVLink pattern ceiling at n=2/4/8,
# geometry sweep at n=8, the real-multi-GPU parity / soak tests, and the driver's bench at N=8
# (f32 -> multicast path, and the bf16 wire).  Everything lands in gpurun_out/.
cd "$(dirname "$0")/.."
O=gpurun_out
./tools/nvlink_ceiling 20 > $O/r02_nvlink_ceiling_n8box.jsonl 2>&1
TUNE_N=${TUNE_N:-8} python tools/tune_r02.py > $O/r02_tune_n8.jsonl 2> $O/r02_tune_n8.err
B2DP_REPORT_SWITCH_ROUNDING=1 python -m pytest tests/test_gpu_multi.py tests/test_gpu_soak.py -q -s \
  -k "(path_b and local) or local_context_distinct or (300_steps and (peer-0 or multicast-0)) or one_host_thread" 2>&1 \
  | grep -E "SWITCH_BF16|passed|failed|FAILED|^E  " | head -40 > $O/r02_pytest_gpu8.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
  bench.py --gpus 8 --steps 20 --warmup 5 > $O/r02_bench_n8.json 2> $O/r02_bench_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 \
  bench.py --gpus 8 --steps 5 --warmup 3 --wire bf16 --no-cpu-baseline > $O/r02_bench_n8_bf16.json 2> $O/r02_bench_n8_bf16.err
cat $O/r02_nvlink_ceiling_n8box.jsonl | grep fused
cat $O/r02_tune_n8.jsonl; tail -2 $O/r02_tune_n8.err
cat $O/r02_pytest_gpu8.log | cut -c1-250
for f in $O/r02_bench_n8.json $O/r02_bench_n8_bf16.json; do python - "$f" <<'PY'
import json, sys
lines = [l for l in open(sys.argv[1]) if l.startswith("{")]
if not lines:
    print(sys.argv[1], "NO JSON LINE")
else:
    d = json.loads(lines[-1])
    print(sys.argv[1], d.get("value"), json.dumps(d.get("parity")), json.dumps(d.get("hot_path")), json.dumps(d.get("roofline"))[:600])
PY
done
tail -3 $O/r02_bench_n8.err $O/r02_bench_n8_bf16.err | cut -c1-300
