#!/bin/bash
# usage (on an N-GPU box): bash tools/n8_one.sh <tag> <N> <workload> — one device-resident bench line of the sharded sort
tag=$1; N=$2; wl=$3
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29841 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --no-e2e --workload $wl > gpurun_out/${tag}_bench_n${N}_$wl.json 2> gpurun_out/${tag}_bench_n${N}_$wl.err
python - gpurun_out/${tag}_bench_n${N}_$wl.json $wl <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
ex = d.get("exchange", {})
print(sys.argv[2], round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms", ex.get("phase_ms"), ex.get("splitter_mode"), ex.get("imbalance"), ex.get("splitter_plan_ms"), (d.get("verified") or {}).get("multiset_checksums_match"), (d.get("verified") or {}).get("records_per_rank"))
PY
