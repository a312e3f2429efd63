#!/usr/bin/env bash
# round-2 ncu evidence: launch lists of two decode steps (bf16, int8 weights, token-ring stage step) and full captures of
# the dominant kernels.  Every command first runs once without ncu.
mkdir -p gpurun_out
for m in decode decode_w8 stage; do
  python tests/prof_step.py $m > gpurun_out/prof_plain_$m.log 2>&1 || { tail -5 gpurun_out/prof_plain_$m.log; exit 1; }
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/launches_$m.csv python tests/prof_step.py $m > gpurun_out/ncu_$m.log 2>&1; echo "launch list $m: $?"
done
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_tcgen05_kernel -c 4 \
    -o gpurun_out/prof_gemm_decode -f python tests/prof_step.py decode > gpurun_out/ncu_full1.log 2>&1; echo "full gemm: $?"
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_w8_kernel -c 4 \
    -o gpurun_out/prof_gemm_w8 -f python tests/prof_step.py decode_w8 > gpurun_out/ncu_full2.log 2>&1; echo "full w8: $?"
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:argmax_kernel|rmsnorm_kernel|cast_kernel" -c 4 \
    -o gpurun_out/prof_stage_edges -f python tests/prof_step.py stage > gpurun_out/ncu_full3.log 2>&1; echo "full edges: $?"
