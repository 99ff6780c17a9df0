#!/usr/bin/env python
"""Time ocsb_hfun_flow_limiter alone on a resident 20000^2 DEM (CUDA events);
run once per OCSB_LIM_ROWS value by the caller."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ocsmesh_b200 import _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
ctx = _lib.default_context(0)
z = (torch.randn(n, n, device="cuda") * 20).contiguous()
old = torch.full((n, n), 700.0, device="cuda")
new = torch.empty_like(old)
win = _lib.make_win((0.0, 0.0, float(n), float(n)), n, n, 0, 0)
def run(o):
    ctx.hfun_flow_limiter(win, z, _lib.OCSB_F32, n, o, n, new, n, 10.0, 10.0,
                          50.0, 1500.0, None, None, 100.0, 1500.0, True)
for o, name in ((old, "old"), (None, "pristine")):
    for _ in range(3): run(o)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.set_timing(True); ctx.reset_timing()
    for _ in range(10): run(o)
    ctx.sync()
    t = ctx.get_timing()
    ms = t["flow_limiter"][0] / t["flow_limiter"][1]
    print(f"rows={os.environ.get('OCSB_LIM_ROWS','default')} {name}: {ms:.4f} ms  "
          f"{(12 if o is not None else 8) * n * n / ms / 1e6:.0f} GB/s", flush=True)
    ctx.set_timing(False)
