mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r1_smi.txt 2>&1
for l in "0 0" "0 1" "1 1" "1 0"; do
  timeout 120 python tests/probe_gemm.py $l >> gpurun_out/r1_probe.log 2>&1
  echo "exit $? for layout $l" >> gpurun_out/r1_probe.log
done
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q --timeout 120 > gpurun_out/r1_gemm.log 2>&1
timeout 300 python -m pytest tests/test_attn_gpu.py -m gpu -q --timeout 120 > gpurun_out/r1_attn.log 2>&1
timeout 600 python -m pytest tests/test_block_gpu.py -m gpu -q --timeout 200 > gpurun_out/r1_block.log 2>&1
tail -5 gpurun_out/r1_probe.log gpurun_out/r1_gemm.log gpurun_out/r1_attn.log gpurun_out/r1_block.log
