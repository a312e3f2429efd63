#!/usr/bin/env bash
# every documented experiment switch must still give correct results: ops + span suites under each of them
# the switches exist only in the experiment build (DLI_EXPERIMENTS=1 distributed_llm_inference_b200/csrc/build.sh)
export DLI_LIB=${DLI_LIB:-$(pwd)/distributed_llm_inference_b200/csrc/libdli_b200_exp.so}
for sw in DLI_NO_PDL DLI_NO_PAIR DLI_NO_PERSISTENT DLI_DEC_DIRECT DLI_DEC_NO_MMA DLI_NO_ROPE_FUSE DLI_ATTN_LEGACY DLI_FA_V1; do
  r=$(env $sw=1 timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_span_gpu.py -q -m gpu -p no:cacheprovider -x 2>&1 | tail -1)
  echo "$sw=1: $r"
done
