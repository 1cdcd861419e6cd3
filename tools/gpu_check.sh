#!/bin/bash
# tools/gpu_check.sh [bench args]: on the GPU box, the GPU test-suite then bench.py; the JSON
# line goes to gpurun_out/bench_last.json with a short digest on stdout.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
timeout 1200 python bench.py "$@" > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err
tail -c 800 gpurun_out/bench_last.err
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_last.json"))
except Exception as e:
    raise SystemExit(f"no bench line: {e}")
def show(k, v):
    r = v.get("roofline", {})
    print(k, "value %.4g" % v.get("value", 0), "ms/step %.4f" % v.get("ms_per_step", 0), "rounds", v.get("rounds"),
          "frac %.3f" % r.get("frac", 0), "kern_ms %.4f" % r.get("avg_launch_ms", 0), "e2e %.4g" % v.get("e2e", {}).get("value", 0),
          v.get("error", ""), v.get("sharded_equals_unsharded", ""), v.get("e2e_equals_device_path", ""))
show("main", d)
for k, v in d.get("workloads", {}).items():
    if "error" in v:
        print(k, v["error"])
    else:
        show(k, v)
PY
