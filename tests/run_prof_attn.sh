mkdir -p gpurun_out
BB=16 timeout 300 python tests/bench_kernels.py > gpurun_out/r7_plain.log 2>&1 && \
BB=16 timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_tc -s 1 -c 1 -o gpurun_out/r7_attn python tests/bench_kernels.py > gpurun_out/r7_ncu.log 2>&1
tail -n 3 gpurun_out/r7_ncu.log
