#!/bin/bash
# usage (on an 8-GPU box): bash tools/n8_quick.sh <tag> — parity of the sharded sort on 8 GPUs (peer mode, incl. the 40 M-per-rank case),
# then the default bench line and the hashed-Zipf workload at N=8 (device-resident value only)
tag=${1:-n8q}
mkdir -p gpurun_out
DIST_CHECK_MODE=peer DIST_CHECK_BIG=40000000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 tests/dist_check.py 2>&1 | grep "dist_check\|rror" > gpurun_out/${tag}_dist_check_n8_peer.log
echo "dist_check N=8 peer: $(grep -c ': ok' gpurun_out/${tag}_dist_check_n8_peer.log) ok, $(grep -c MISMATCH gpurun_out/${tag}_dist_check_n8_peer.log) mismatch"
port=29820
for wl in u64_1b_uniform u64_1b_zipf_hashed; do
port=$((port + 1))
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-e2e --workload $wl > gpurun_out/${tag}_bench_n8_$wl.json 2> gpurun_out/${tag}_bench_n8_$wl.err
python - gpurun_out/${tag}_bench_n8_$wl.json $wl <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
ex = d.get("exchange", {})
print(sys.argv[2], round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms", ex.get("phase_ms"), ex.get("splitter_mode"), ex.get("imbalance"), (d.get("verified") or {}).get("multiset_checksums_match"), d.get("route"))
PY
done
