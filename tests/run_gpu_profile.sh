#!/usr/bin/env bash
mkdir -p gpurun_out
run() { name=$1; shift; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; rc=$?; echo "$name exit $rc"; tail -n ${TAILN:-8} gpurun_out/$name.log; return $rc; }
run zz_alone python -m pytest tests/test_zz_no_alloc.py -q -m gpu --no-header -p no:cacheprovider
run all_gpu python -m pytest tests -x -q -m gpu --no-header -p no:cacheprovider
KN='regex:gemm_tcgen05|rmsnorm_kernel|rope_kv|decode_attn|decode_combine|prefill_attn|silu_mul|cast_kernel'
CMD="python bench.py --steps 2 --warmup 3 --cpu-budget 1"
run plain $CMD && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KN" -s 289 -c 706 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"; tail -n 3 gpurun_out/ncu_list.log
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 256 -c 4 -o gpurun_out/prof_gemm_decode $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -n 3 gpurun_out/ncu_full.log
ncu --set full --clock-control none --import-source on -k regex:decode_attn_kernel -s 32 -c 1 -o gpurun_out/prof_attn_decode $CMD > gpurun_out/ncu_full2.log 2>&1
echo "ncu full2 exit $?"; ls -la gpurun_out
