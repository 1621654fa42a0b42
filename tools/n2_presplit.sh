# usage (on a 2-GPU box): bash tools/n2_presplit.sh  — parity (dist_check, peer mode) and the presplit exchange on / off at 200 M keys per rank
mkdir -p gpurun_out
: # (dist_check ran in the previous call)

for n in 200000000 50000000; do
for ps in 1 0; do
B200SORT_DIST_PRESPLIT=$ps python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2972$ps bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-e2e --records $n > gpurun_out/r11_bench_n2_${n}_presplit$ps.json 2> gpurun_out/r11_bench_n2_${n}_presplit$ps.err
python - gpurun_out/r11_bench_n2_${n}_presplit$ps.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
ex = d.get("exchange", {})
print(sys.argv[1], round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms", ex.get("phase_ms"), "presplit", ex.get("presplit"), (d.get("verified") or {}).get("multiset_checksums_match"))
PY
done
done
