#!/bin/bash
# Round-2 measurement session on a 4-GPU box (gpurun --gpus 4).
cd "$(dirname "$0")/.."
O=gpurun_out
if [ -z "$SKIP_TUNE" ]; then TUNE_N=4 python tools/tune_r02.py > $O/r02_tune_n4.jsonl 2> $O/r02_tune_n4.err; fi
python -m pytest tests/test_gpu_multi.py tests/test_gpu_soak.py tests/test_gpu_examples.py -q ${PYTEST_K:+-k "$PYTEST_K"} 2>&1 | grep -E "passed|failed|FAILED|^E  " | head -30 > $O/r02_pytest_gpu4.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus 4 --steps 20 --warmup 5 > $O/r02_bench_n4.json 2> $O/r02_bench_n4.err
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke4.log 2>&1
cat $O/r02_tune_n4.jsonl; tail -2 $O/r02_tune_n4.err; cat $O/r02_pytest_gpu4.log | cut -c1-250; tail -2 $O/r02_smoke4.log
python - <<'PY'
import json
lines = [l for l in open("gpurun_out/r02_bench_n4.json") if l.startswith("{")]
d = json.loads(lines[-1]) if lines else {}
print(d.get("value"), json.dumps(d.get("parity")), json.dumps(d.get("hot_path")), json.dumps(d.get("roofline"))[:700], json.dumps(d.get("cpu_baseline"))[:200])
PY
tail -3 $O/r02_bench_n4.err | cut -c1-300
