# bench line + ncu launch list of the same command + ncu --set full of the column-sum kernel (round-1 closing capture)
mkdir -p gpurun_out
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?" >> gpurun_out/k_bench.err
cat gpurun_out/k_bench.json; tail -2 gpurun_out/k_bench.err
timeout 150 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/k_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 5000 --csv --log-file gpurun_out/k_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/k_ncu.log 2>&1
tail -2 gpurun_out/k_ncu.log
LAYERS=1 PLAIN=1 DROPOUT=0.1 GXA_PDL=0 timeout 200 ncu --set full --clock-control none --import-source on -k regex:colsum_wide -s 6 -c 3 -o gpurun_out/k_colsum python tools/step_profile.py > gpurun_out/k_ncu2.log 2>&1
tail -2 gpurun_out/k_ncu2.log
