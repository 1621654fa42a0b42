# usage (on a 2-GPU box): bash tools/n2_check.sh <tag> — parity of the sharded sort (tests/dist_check.py, peer mode, incl. the 40 M-per-rank case) and the default bench line at N=2
tag=${1:-n2}
mkdir -p gpurun_out
DIST_CHECK_MODE=peer DIST_CHECK_BIG=40000000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 tests/dist_check.py 2>&1 | grep "dist_check\|rror" > gpurun_out/${tag}_dist_check_n2_peer.log
echo "dist_check: $(grep -c ': ok' gpurun_out/${tag}_dist_check_n2_peer.log) ok, $(grep -c MISMATCH gpurun_out/${tag}_dist_check_n2_peer.log) mismatch"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/${tag}_bench_n2.json 2> gpurun_out/${tag}_bench_n2.err
python - gpurun_out/${tag}_bench_n2.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
ex = d.get("exchange", {})
print(round(d["value"], 2), "Gkeys/s", round(d["ms_per_step"], 2), "ms", ex.get("phase_ms"), (d.get("verified") or {}).get("multiset_checksums_match"))
PY
