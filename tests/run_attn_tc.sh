mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_attn_gpu.py -m gpu -q --timeout 120 -x > gpurun_out/r6_attn.log 2>&1
timeout 600 python -m pytest tests/test_block_gpu.py -m gpu -q --timeout 200 > gpurun_out/r6_block.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r6_bench.json 2> gpurun_out/r6_bench.err
tail -n 25 gpurun_out/r6_attn.log; tail -n 5 gpurun_out/r6_block.log; python -c "
import json; d=json.load(open('gpurun_out/r6_bench.json')); print(d['ms_per_step'], d['value'], d['step_roofline']['frac'])"; tail -n 3 gpurun_out/r6_bench.err
