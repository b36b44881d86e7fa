mkdir -p gpurun_out
DROPOUT=0.1 timeout 120 python tools/step_profile.py > gpurun_out/h_step64.txt 2>&1
DROPOUT=0.1 timeout 100 python tools/step_profile.py whisper-medium-c3 8 > gpurun_out/h_step8.txt 2>&1
tail -18 gpurun_out/h_step64.txt; tail -18 gpurun_out/h_step8.txt
