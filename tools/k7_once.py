#!/usr/bin/env python
"""Three launches of one K7 variant on the C2 ring (for ncu captures):
   K7_VARIANT=hinted|stream|warp|block (default: klm_dev_window_stats), K7_DIST=uniform|quantised."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
dev = pkg.Dev()
s, t = int(os.environ.get("K7_SERIES", 26000)), int(os.environ.get("K7_WINDOW", 3600))
pitch = (t + 3) & ~3
g = torch.Generator(device="cuda").manual_seed(1234)
ring = torch.rand((s, pitch), device="cuda", generator=g) * 100
if os.environ.get("K7_DIST") == "quantised":
    ring = torch.round(ring / 5) * 5
ring = ring.float().contiguous()
seen = torch.full((s,), t, dtype=torch.int32, device="cuda")
o = [torch.empty(s, dtype=torch.float32, device="cuda") for _ in range(3)]
om = torch.empty(s, dtype=torch.float64, device="cuda")
variant = os.environ.get("K7_VARIANT") or None
if variant == "hinted":   # streaming: a fresh sample per series between the calls, brackets carried in `state`
    state = torch.full((s, 2), float("nan"), dtype=torch.float32, device="cuda")
    k7 = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], state=state)
    for i in range(4):
        dev.ring_ingest(ring, seen, t, (torch.rand(s, device="cuda", generator=g) * 100).float())
        k7()
else:
    k7 = dev.prebind("window_stats", ring, seen, t, o[0], o[1], om, o[2], variant=variant)
    for _ in range(3):
        k7()
torch.cuda.synchronize()
print("ok", float(o[2][0]))
