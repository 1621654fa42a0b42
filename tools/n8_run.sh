#!/bin/bash
# The 8-GPU evidence run of a round (on the GPU box): multi-GPU parity logs for N = 2, 4, 8 x {peer, nccl, ce}
# (tests/dist_check.py against the oracle, plus a >= 32 Mi-records-per-rank case checked on the devices), the
# single-process entry points on 1/2/4/8 GPUs, then bench lines: N=8 all configs[4] workloads, N=4 default.
tag=${1:-n8}
mkdir -p gpurun_out
port=29700
for w in 8 4 2; do
  for mode in peer nccl ce; do
    port=$((port + 1))
    big=""
    if [[ $mode == peer || $w == 8 ]]; then big=40000000; fi
    DIST_CHECK_MODE=$mode DIST_CHECK_BIG=$big timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w \
      --master-addr 127.0.0.1 --master-port $port tests/dist_check.py 2>&1 | grep "dist_check\|rror" > gpurun_out/${tag}_dist_check_n${w}_${mode}.log
    echo "dist_check N=$w $mode: $(grep -c ': ok' gpurun_out/${tag}_dist_check_n${w}_${mode}.log) ok, $(grep -c MISMATCH gpurun_out/${tag}_dist_check_n${w}_${mode}.log) mismatch"
  done
done
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -2 | tee gpurun_out/${tag}_pytest_multi.log
port=$((port + 1))
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
  bench.py --gpus 8 --steps 5 --warmup 3 --all-configs > gpurun_out/${tag}_bench_n8_all_configs.jsonl 2> gpurun_out/${tag}_bench_n8.err
tail -c 600 gpurun_out/${tag}_bench_n8.err
port=$((port + 1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $port \
  bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/${tag}_bench_n4.json 2> gpurun_out/${tag}_bench_n4.err
python - gpurun_out/${tag}_bench_n8_all_configs.jsonl gpurun_out/${tag}_bench_n4.json <<'PY'
import json, sys
for f in sys.argv[1:]:
    for ln in open(f):
        if not ln.startswith("{"):
            continue
        d = json.loads(ln)
        ex = d.get("exchange", {})
        print(d["n_gpus"], d["config"]["workload"][:22], round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms; e2e",
              (d.get("e2e") or {}).get("value"), ex.get("phase_ms"), ex.get("splitter_mode"), round(ex.get("imbalance", 0), 3), d.get("verified"))
PY
