# ncu --set full of the HBM-bound row-wise kernels inside a real one-layer step at config-3 shapes (B = 64, dropout 0.1)
mkdir -p gpurun_out
export LAYERS=1 PLAIN=1 DROPOUT=0.1 GXA_PDL=0
timeout 120 python tools/step_profile.py > gpurun_out/g_plain.log 2>&1 || { tail -5 gpurun_out/g_plain.log; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on \
  -k 'regex:ln_fwd|ln_bwd|gate_bwd|colsum|attn_delta|dq_finalize|cast_segs' -s 26 -c 14 -o gpurun_out/g_rowwise \
  python tools/step_profile.py > gpurun_out/g_ncu.log 2>&1
tail -n 3 gpurun_out/g_ncu.log; ls -la gpurun_out/g_rowwise.ncu-rep
