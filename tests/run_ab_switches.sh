# A/B of the experiment switches that are off by default; every line is a full bench.py run
# (device-timed ms/step), so box-to-box clock differences do not enter the comparison.  Build the variant library first,
# on the CPU side:   GXA_BUILD_VARIANT=dpx GXA_NVCC_EXTRA=-DGXA_DROP_DPX python -m mocov2_whisper_flamingo_b200.build
mkdir -p gpurun_out
run() {   # label, env...
  label=$1; shift
  env "$@" timeout 150 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err
  python - "$label" <<'PY'
import json, sys
try:
    d = json.load(open(f"gpurun_out/ab_{sys.argv[1]}.json"))
    print(f"{sys.argv[1]:16s} {d['ms_per_step']:8.2f} ms/step  e2e {d['e2e']['ms_per_step']:8.2f}  sm {d['clocks']['sm_mhz']} MHz")
except Exception as e:
    print(sys.argv[1], "failed:", e)
PY
}
run base X=0
run wgrad_stages5 GXA_WGRAD_STAGES=5
run wgrad_stages4 GXA_WGRAD_STAGES=4
[ -f mocov2_whisper_flamingo_b200/libgxa_b200.dpx.so ] && run drop_dpx GXA_LIB=$PWD/mocov2_whisper_flamingo_b200/libgxa_b200.dpx.so
run base_again X=0
