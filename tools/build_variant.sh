#!/bin/bash
# tools/build_variant.sh NAME "EXTRA_NVCC_FLAGS": builds k8s-llm-monitor_b200/libklm_NAME.so from
# the same sources with extra flags (ablation and tuning experiments; select it at run time
# with KLM_LIB_NAME=libklm_NAME.so). Objects go to a scratch directory, the normal build
# is untouched.
set -e
name=$1; extra=$2
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/k8s-llm-monitor_b200/csrc
obj=$(mktemp -d)
pids=
for f in abi schedule aggregate ring window_stats_warp window_stats_stream noalloc predicates topk; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo \
    -Xcompiler -fPIC,-Wall,-fvisibility=default --expt-relaxed-constexpr $extra \
    -c $src/$f.cu -o $obj/$f.o &
  pids="$pids $!"
done
for p in $pids; do wait $p; done   # set -e: a failed compile stops here
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared \
  -o $root/k8s-llm-monitor_b200/libklm_$name.so $obj/*.o -cudart shared
rm -rf $obj
echo built libklm_$name.so
