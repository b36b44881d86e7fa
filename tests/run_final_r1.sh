# Round-1 closing GPU call (8 GPU-minutes left): parity of the tests that exercise the rewritten column-sum kernel,
# the bench line, the stock-torch eager GPU baseline and the in-step kernel breakdown.
mkdir -p gpurun_out
timeout 170 python -m pytest tests/test_block_gpu.py -m gpu -x -q > gpurun_out/f_tests.log 2>&1; echo "rc=$?" >> gpurun_out/f_tests.log
tail -3 gpurun_out/f_tests.log
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "rc=$?" >> gpurun_out/f_bench.err
cat gpurun_out/f_bench.json; tail -2 gpurun_out/f_bench.err
timeout 120 python tests/bench_torch_eager_gpu.py > gpurun_out/f_eager.json 2> gpurun_out/f_eager.err; echo "rc=$?" >> gpurun_out/f_eager.err
cat gpurun_out/f_eager.json; tail -2 gpurun_out/f_eager.err
DROPOUT=0.1 timeout 120 python tools/step_profile.py > gpurun_out/f_step.txt 2>&1; echo "rc=$?" >> gpurun_out/f_step.txt
head -30 gpurun_out/f_step.txt
GXA_COLSUM_NARROW=1 timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/f_bench_narrow.json 2> gpurun_out/f_bench_narrow.err
cat gpurun_out/f_bench_narrow.json
