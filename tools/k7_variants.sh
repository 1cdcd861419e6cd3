#!/bin/bash
# tools/k7_variants.sh LIB "V1 V2 ...": times the hinted K7 of a tuning build for each value of
# KLM_WSS_VARIANT (tools/k7_tune.py worker), after checking the hinted ring tests with it.
lib=$1; shift
mkdir -p gpurun_out
for v in $1; do
  echo "== variant $v"
  KLM_LIB_NAME=$lib KLM_WSS_VARIANT=$v timeout 600 python -m pytest tests/test_gpu_ring.py -x -q -m gpu -k "hint or state or stream or bracket" 2>&1 | tail -2
  KLM_LIB_NAME=$lib KLM_WSS_VARIANT=$v K7_WORKER=1 timeout 300 python tools/k7_tune.py | tee -a gpurun_out/k7_variants.jsonl
done
