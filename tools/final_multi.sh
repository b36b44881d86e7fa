#!/bin/bash
# Multi-GPU evidence: bash tools/final_multi.sh <N> <tag> [c4]   (run through gpurun --gpus N)
N=$1; tag=${2:-r2z}; out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
set -x
TIMELINE=0 run 29541 tools/dp_profile.py > $out/${tag}_dp_profile_n$N.txt 2>&1
run 29542 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-gpu-baseline > $out/${tag}_scale_n$N.json 2> $out/${tag}_scale_n$N.err
if [ "$3" == "c4" ]; then
  run 29544 bench.py --gpus $N --workload whisper-large-c4 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline > $out/${tag}_c4_n${N}_contiguous.json 2> $out/${tag}_c4_contig.err
  run 29545 bench.py --gpus $N --workload whisper-large-c4 --shards balanced --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline > $out/${tag}_c4_n${N}_balanced.json 2> $out/${tag}_c4_bal.err
fi
grep "N=$N" $out/${tag}_dp_profile_n$N.txt
for f in $out/${tag}_scale_n$N*.json $out/${tag}_c4_n${N}_*.json; do [ -f $f ] && python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["clocks"]["sm_mhz"], d.get("rank_imbalance_max_over_mean"), d.get("dp_grad_check", {}).get("max_rel_err"), d.get("dp_grad_check", {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "ERR", e)
PY
done
