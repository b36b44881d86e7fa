mkdir -p gpurun_out
timeout 300 python tests/prof_kernels.py gemm > gpurun_out/r5_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 7 -o gpurun_out/r5_gemm python tests/prof_kernels.py gemm > gpurun_out/r5_ncu.log 2>&1
tail -n 3 gpurun_out/r5_ncu.log
