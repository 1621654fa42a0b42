#!/bin/bash
# Multi-GPU runs on one box.  usage (on the GPU box): bash tools/scale_run.sh <tag> "<N list>" [tests] [zipf]
#   -> gpurun_out/<tag>_bench_n<N>.json per N; "tests" runs the distributed parity tests first,
#      "zipf" adds the skewed workload at the largest N.
tag=${1:-scale}
ns=${2:-"8 4 2"}
mkdir -p gpurun_out
if [[ " $* " == *" tests "* ]]; then python -m pytest tests -m gpu -x -q -k "dist or sharded or multi" 2>&1 | tail -2; fi
port=29600
show() {
python - "$1" "$2" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
ex = d.get("exchange", {})
print(sys.argv[2], round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms; e2e", (d.get("e2e") or {}).get("value"),
      ex.get("phase_ms"), ex.get("splitter_mode"), ex.get("imbalance"))
PY
}
for n in $ns; do
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 5 --warmup 3 --no-cpu > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err
  show gpurun_out/${tag}_bench_n$n.json $n
done
if [[ " $* " == *" zipf "* ]]; then
  n=${ns%% *}
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 5 --warmup 3 --no-cpu --no-e2e --workload u64_1b_zipf_hashed \
    > gpurun_out/${tag}_bench_n${n}_zipf_hashed.json 2> gpurun_out/${tag}_bench_n${n}_zipf_hashed.err
  show gpurun_out/${tag}_bench_n${n}_zipf_hashed.json "zipf_hashed $n"
fi
