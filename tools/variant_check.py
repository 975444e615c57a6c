"""Parity + timing of one library build (BRICKS_LIB) for selected stencils: one 512^3 subdomain
(66^3 bricks incl. ghosts).  Parity: one sweep against the naive verification twin of the SAME
library (max relative error) and a CRC of the output so two builds can be compared bit for bit.
Timing: 20 ping-pong sweeps between CUDA events.  Experiment tooling; results are throw-away."""
import os, sys, zlib
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from bricklib_b200 import capi, synth


def main():
    which = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["125"])]
    tag = os.environ.get("VARIANT", "product")
    ctx = capi.Context(0)
    dec = capi.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    f = capi.uniform(nb * 512, seed=3)
    d_a, d_b, d_c = ctx.upload(f), ctx.malloc(nb * 4096), ctx.malloc(nb * 4096)
    plan = capi.Plan(ctx, dec.grid, dec.adj)
    for w in which:
        cf = synth.coefficients()[:27] if w in (0, 1, 27) else None
        ctx.memset(d_b, 0, nb * 4096)
        ctx.memset(d_c, 0, nb * 4096)
        plan.stencil(w, d_a, d_b, coeff=cf)
        plan.stencil_naive(w, d_a, d_c, coeff=cf)
        ctx.sync()
        got = ctx.download(d_b, nb * 512)
        exp = ctx.download(d_c, nb * 512)
        err = float(np.max(np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)))
        crc = zlib.crc32(got.tobytes())
        for _ in range(2):
            plan.stencil(w, d_a, d_b, coeff=cf); plan.stencil(w, d_b, d_c, coeff=cf)
        ctx.timer_start(0)
        for _ in range(10):
            plan.stencil(w, d_a, d_b, coeff=cf); plan.stencil(w, d_b, d_c, coeff=cf)
        ctx.timer_stop(0)
        ms = ctx.timer_ms(0) / 20
        print("VCHECK variant=%s maxcol=%s stencil=%d maxrel=%.2e crc=%08x ms=%.4f GB/s=%.0f" % (
            tag, os.environ.get("BRICKS_MAXCOL", "-"), w, err, crc, ms, 66**3 * 512 * 16 / ms * 1e-6), flush=True)


main()
