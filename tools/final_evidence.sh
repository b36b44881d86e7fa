#!/bin/bash
# One-GPU evidence pass for profiles/ (run on the GPU box through gpurun; every ncu capture follows a plain run of the
# same command that exited 0).  Writes into gpurun_out/<tag>_*; copy what is to be judged into profiles/.
#   bash tools/final_evidence.sh r2z [quick]
tag=${1:-r2z}
out=gpurun_out
mkdir -p $out
set -x
python -m pytest tests -x -q -m gpu > $out/${tag}_tests.log 2>&1; echo "tests rc=$?"
python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_c3.json 2> $out/${tag}_bench_c3.err; echo "c3 rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_c3_reference_arm.json 2> $out/${tag}_ref.err; echo "ref rc=$?"
python bench.py --workload whisper-small-c2 --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_bench_c2.json 2> $out/${tag}_c2.err; echo "c2 rc=$?"
python bench.py --workload whisper-large-c4 --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_c4.json 2> $out/${tag}_c4.err; echo "c4 rc=$?"
DROPOUT=0.1 python tools/step_profile.py whisper-medium-c3 64 > $out/${tag}_step_breakdown_b64.txt 2>&1
DROPOUT=0.1 python tools/step_profile.py whisper-medium-c3 8 > $out/${tag}_step_breakdown_b8.txt 2>&1
python tools/bench_decode.py > $out/${tag}_decode_c5.json 2> $out/${tag}_decode.err
python tools/prof_attn.py 64 0.1 > $out/${tag}_attn_timing.txt 2>&1
python tools/bench_glue.py > $out/${tag}_glue.txt 2>&1
if [ "$2" != "quick" ]; then
  # launch list of the bench command (3 warm-up + 1 timed step + the e2e steps): per-kernel SHARES, cold-cache / serialised
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-baseline > $out/${tag}_plain_for_ncu.json 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/${tag}_launches.csv \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-baseline > $out/${tag}_ncu_launches.log 2>&1
  python tools/ncu_summary.py launches $out/${tag}_launches.csv > $out/${tag}_launches_bench_c3.txt 2>&1
  rm -f $out/${tag}_launches.csv
  # attention kernels with training dropout under ncu --set full (after a plain run of the same command above)
  ncu --set full --clock-control none --import-source on -k regex:attn_fwd_tc2 -c 1 -o $out/${tag}_attn_fwd \
      python tools/prof_attn.py 64 0.1 > $out/${tag}_ncu_attn_fwd.log 2>&1
  python tools/ncu_summary.py rep $out/${tag}_attn_fwd.ncu-rep > $out/${tag}_ncu_attn_fwd.txt 2>&1
  ncu -i $out/${tag}_attn_fwd.ncu-rep --page source --csv > $out/${tag}_attn_fwd_src.csv 2>/dev/null
  python tools/ncu_opmix.py $out/${tag}_attn_fwd_src.csv $((64*16*448*750)) >> $out/${tag}_ncu_attn_fwd.txt 2>&1
  rm -f $out/${tag}_attn_fwd_src.csv
fi
ls -la $out | tail -30
