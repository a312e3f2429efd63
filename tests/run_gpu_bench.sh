#!/usr/bin/env bash
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
run ops python -m pytest tests/test_ops_gpu.py -q -m gpu --no-header -p no:cacheprovider --maxfail=10
run span python -m pytest tests/test_span_gpu.py -q -m gpu --no-header -p no:cacheprovider -s --maxfail=10
TAILN=30 run bench python bench.py --steps 32 --warmup 4 --cpu-budget 8
run bench_ref python bench.py --impl reference --steps 2 --warmup 1
