#!/usr/bin/env bash
# ncu captures for profiles/ (run under gpurun, one GPU).  Every ncu run repeats a command that has just exited 0.
mkdir -p gpurun_out
KN='regex:gemm_tcgen05|rmsnorm_kernel|rope_kv|decode_attn|decode_combine|prefill_attn|silu_mul|cast_kernel|kv_evict'
CMD="python bench.py --steps 2 --warmup 3 --cpu-budget 0"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
# launch list: two decode steps (225 launches each: 7 per layer + the exit cast) after prefill (225), 4 warm prefills,
# graph warm-up and 3 warm-up steps
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KN" -s 2025 -c 450 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
# launch list of one warm prefill (225 launches)
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KN" -s 225 -c 225 --csv --log-file gpurun_out/launches_prefill.csv $CMD > gpurun_out/ncu_list2.log 2>&1
echo "ncu list prefill exit $?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05_kernel -s 512 -c 4 -o gpurun_out/prof_gemm_decode $CMD > gpurun_out/ncu_full1.log 2>&1
echo "ncu gemm decode exit $?"
ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 128 -c 4 -o gpurun_out/prof_gemm_prefill $CMD > gpurun_out/ncu_full2.log 2>&1
echo "ncu gemm prefill exit $?"
ncu --set full --clock-control none --import-source on -k regex:decode_attn_kernel -s 40 -c 1 -o gpurun_out/prof_attn_decode $CMD > gpurun_out/ncu_full3.log 2>&1
echo "ncu attn decode exit $?"
ncu --set full --clock-control none --import-source on -k regex:prefill_attn_tc -s 33 -c 1 -o gpurun_out/prof_attn_prefill_tc $CMD > gpurun_out/ncu_full4.log 2>&1
echo "ncu attn prefill exit $?"
# grouped-query decode attention: one 70B-shaped stage (10 layers), batch 32
CMD70="python bench.py --steps 2 --warmup 3 --cpu-budget 0 --model 70b --layers 10 --batch 32"
$CMD70 > gpurun_out/plain70.log 2>&1 || { echo "plain 70b run failed"; tail -5 gpurun_out/plain70.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:decode_attn_gqa -s 25 -c 1 -o gpurun_out/prof_attn_decode_gqa $CMD70 > gpurun_out/ncu_full5.log 2>&1
echo "ncu attn decode gqa exit $?"
# long-chunk prefill attention: one 70B-shaped stage, 1 x 8192 tokens
CMD8K="python bench.py --steps 2 --warmup 3 --cpu-budget 0 --model 70b --layers 10 --batch 1 --prompt 8192"
$CMD8K > gpurun_out/plain8k.log 2>&1 || { echo "plain 8k run failed"; tail -5 gpurun_out/plain8k.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:prefill_attn_tc2 -s 11 -c 1 -o gpurun_out/prof_attn_prefill_tc2 $CMD8K > gpurun_out/ncu_full6.log 2>&1
echo "ncu attn prefill tc2 exit $?"
ls -la gpurun_out | tail -18
