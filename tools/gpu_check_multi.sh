#!/bin/bash
# tools/gpu_check_multi.sh N [bench args]: on a box with N GPUs, the sharding tests (multi-device
# handle included), then bench.py under torchrun on N ranks.
n=$1; shift
mkdir -p gpurun_out
nvidia-smi topo -m 2>&1 | head -12 > gpurun_out/topo_n$n.txt
timeout 600 python -m pytest tests/test_gpu_shard.py tests/test_gpu_host.py -x -q -m gpu 2>&1 | tail -6
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $n "$@" > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
echo "bench rc=$?"
grep -v "^W1\|^\*\*\*\|OMP_NUM" gpurun_out/bench_n$n.err | tail -c 1500
python - $n <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/bench_n{n}.json") if l.startswith("{")][-1])
except Exception as e:
    raise SystemExit(f"no bench line: {e}")
def show(k, v):
    r = v.get("roofline", {})
    print(k, "value %.4g" % v.get("value", 0), "ms/step %.4f" % v.get("ms_per_step", 0), "rounds", v.get("rounds"),
          "frac %.3f" % r.get("frac", 0), "kern_ms %.4f" % r.get("avg_launch_ms", 0), "e2e %.4g" % v.get("e2e", {}).get("value", 0),
          v.get("sharded_equals_unsharded", ""), v.get("e2e_equals_device_path", ""))
show("main", d)
for k, v in d.get("workloads", {}).items():
    if "error" in v:
        print(k, v["error"])
    else:
        show(k, v)
PY
