# Round-1 closing GPU calls: parity of the tests that exercise the rewritten column-sum kernel, the bench line,
# and the in-step kernel breakdown.  (tests/bench_torch_eager_gpu.py gives the stock-torch eager GPU baseline)
mkdir -p gpurun_out
timeout 170 python -m pytest tests/test_block_gpu.py tests/test_avnet_gpu.py tests/test_avglue.py -m gpu -x -q > gpurun_out/f_tests.log 2>&1; echo "rc=$?" >> gpurun_out/f_tests.log
tail -3 gpurun_out/f_tests.log
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "rc=$?" >> gpurun_out/f_bench.err
cat gpurun_out/f_bench.json; tail -2 gpurun_out/f_bench.err
DROPOUT=0.1 timeout 120 python tools/step_profile.py > gpurun_out/f_step.txt 2>&1; echo "rc=$?" >> gpurun_out/f_step.txt
head -30 gpurun_out/f_step.txt
