mkdir -p gpurun_out
rm -f gpurun_out/r4_*.log
for l in "0 0" "0 1" "1 1" "1 0"; do
  timeout 120 python tests/probe_gemm2.py $l >> gpurun_out/r4_probe.log 2>&1
  echo "exit $? for layout $l" >> gpurun_out/r4_probe.log
done
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q --timeout 120 > gpurun_out/r4_gemm.log 2>&1
timeout 600 python -m pytest tests/test_block_gpu.py tests/test_attn_gpu.py -m gpu -q --timeout 200 > gpurun_out/r4_block.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r4_bench.json 2> gpurun_out/r4_bench.err
cat gpurun_out/r4_probe.log; tail -n 5 gpurun_out/r4_gemm.log; tail -n 5 gpurun_out/r4_block.log; cat gpurun_out/r4_bench.json; tail -n 3 gpurun_out/r4_bench.err
