"""Time one full-grid sweep (66^3 bricks) of selected stencils; env vars pick kernel variants."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from bricklib_b200 import capi

def main():
    which = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["7", "13", "25", "125"])]
    ctx = capi.Context(0)
    dec = capi.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    f = capi.uniform(nb * 512, seed=3)
    d_a, d_b = ctx.upload(f), ctx.malloc(nb * 4096)
    ctx.memset(d_b, 0, nb * 4096)
    plan = capi.Plan(ctx, dec.grid, dec.adj)
    for w in which:
        ctx.h2d(d_a, f); ctx.sync()
        for _ in range(2):
            plan.stencil(w, d_a, d_b); plan.stencil(w, d_b, d_a)
        ctx.timer_start(0)
        for _ in range(10):
            plan.stencil(w, d_a, d_b); plan.stencil(w, d_b, d_a)
        ctx.timer_stop(0)
        ms = ctx.timer_ms(0) / 20
        print("KTIME stencil=%d env=%s ms=%.4f GB/s=%.0f" % (w, {k: v for k, v in os.environ.items() if k.startswith("BRICKS_")}, ms, 66**3*512*16/ms*1e-6), flush=True)
main()
