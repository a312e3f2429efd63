#!/usr/bin/env bash
# Run the GPU test groups in separate processes (a trapped kernel poisons its CUDA context) and keep the logs.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?"; tail -n 25 gpurun_out/$name.log; }
run ops_misc python -m pytest tests/test_ops_gpu.py -q -m gpu -k "not gemm" -x --no-header -p no:cacheprovider
run ops_gemm python -m pytest tests/test_ops_gpu.py -q -m gpu -k "gemm" --no-header -p no:cacheprovider --maxfail=12
run span python -m pytest tests/test_span_gpu.py -q -m gpu --no-header -p no:cacheprovider -s --maxfail=6
run smoke python -c "import __graft_entry__ as g; g.smoke()"
