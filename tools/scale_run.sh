#!/bin/bash
# Multi-GPU run on one box: distributed parity tests, then bench.py at N = 8, 4, 2 (+ the skewed workload at N = 8).
# usage (on the GPU box): bash tools/scale_run.sh <tag>     -> gpurun_out/<tag>_bench_n*.json
tag=${1:-scale}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "dist or sharded or multi" 2>&1 | tail -2
port=29600
for n in 8 4 2; do
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 5 --warmup 3 --no-cpu > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/${tag}_bench_n$n.json").read().strip().splitlines()[-1])
print($n, round(d["value"], 2), round(d["ms_per_step"], 2), d.get("e2e", {}).get("value"), d.get("exchange", {}).get("phase_ms"))
PY
done
port=$((port + 1))
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
  bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-e2e --workload u64_1b_zipf_hashed > gpurun_out/${tag}_bench_n8_zipf.json 2> gpurun_out/${tag}_bench_n8_zipf.err
python - <<PY
import json
d = json.loads(open("gpurun_out/${tag}_bench_n8_zipf.json").read().strip().splitlines()[-1])
print("zipf 8", round(d["value"], 2), round(d["ms_per_step"], 2), d.get("exchange", {}).get("phase_ms"), d.get("exchange", {}).get("splitter_mode"), d.get("exchange", {}).get("imbalance"))
PY
