mkdir -p gpurun_out
python __graft_entry__.py --smoke > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?" >> gpurun_out/r2_bench.err
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2_ncu.log 2>&1
tail -3 gpurun_out/r2_smoke.log; cat gpurun_out/r2_bench.json; tail -5 gpurun_out/r2_bench.err
