"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py: average
duration per kernel of libklm.so and its share of one step (the sum over our kernels).
usage: python tools/launch_list_summary.py launches.csv "<command that was profiled>" """
import csv
import sys
from collections import OrderedDict


def main(path, cmd):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r["Metric Unit"]
            us = v / 1e3 if unit in ("ns", "nsecond") else v * (1e3 if unit == "ms" else 1.0)
            rows.append((r["Kernel Name"], us))
    ours = OrderedDict()
    for name, us in rows:
        if "klm::" not in name:
            continue
        short = name.split("klm::")[1].split("(")[0].split("<")[0]
        ours.setdefault(short, []).append(us)
    setup = {k: v for k, v in ours.items() if k in ("ring_load_kernel",)}  # state load, not a step
    for k in setup:
        del ours[k]
    steps = max(len(v) for v in ours.values()) if ours else 0
    total = sum(sum(v) / len(v) * (len(v) / steps) for v in ours.values()) if steps else 0.0
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none, {cmd}")
    print("# per-launch times are cold-cache and serialised: compare shares, not absolutes")
    print(f"steps found (warm-up + timed + e2e legs): {steps}")
    for k, v in ours.items():
        avg = sum(v) / len(v)
        per_step = len(v) / steps
        print(f"{k:28s} avg {avg:8.2f} us  x{per_step:4.1f}/step  share of step {100 * avg * per_step / total:5.1f}%")
    print(f"step total (sum of kernels) avg {total:.2f} us")
    for k, v in setup.items():
        print(f"(setup, outside the steps: {k} x{len(v)}, avg {sum(v) / len(v):.2f} us)")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "python bench.py")
