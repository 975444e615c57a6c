"""GPU bring-up: hardware probes, TMA swizzle check, small-domain parity of every stencil
against the oracle, and first timings at 512^3.  Run on the B200 box:
    python tools/gpu_bringup.py [--quick] > gpurun_out/bringup.log
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bricklib_b200 import capi  # noqa: E402
from oracle import oracle as O  # noqa: E402

STENCILS = (0, 7, 13, 25, 125)


def relerr(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))


def probes(ctx, out):
    out["dfma_tflops"] = capi.probe_dfma(ctx, 20000)
    out["copy_gbs_2GiB"] = capi.probe_copy(ctx, 2 << 30, 10)
    out["copy_gbs_32MiB"] = capi.probe_copy(ctx, 32 << 20, 50)
    out["read_gbs_2GiB"] = capi.probe_read(ctx, 2 << 30, 10)
    out["read_gbs_32MiB"] = capi.probe_read(ctx, 32 << 20, 50)
    out["read_gbs_8MiB"] = capi.probe_read(ctx, 8 << 20, 100)
    out["mio_lds_only"] = capi.probe_mio(ctx, 1)
    out["mio_shfl_only"] = capi.probe_mio(ctx, 2)
    out["mio_both"] = capi.probe_mio(ctx, 3)
    print("PROBES", json.dumps(out), flush=True)


def swizzle_check(ctx):
    nb = 4
    host = np.arange(nb * 512, dtype=np.float64)
    d = ctx.upload(host)
    img = capi.probe_tma_image(ctx, d, nb, 512, 2)
    ctx.free(d)
    base = 2 * 512
    exp = np.zeros(512)
    for k in range(8):
        for j in range(8):
            for chunk in range(4):
                phys = chunk ^ ((j >> 1) & 3)
                for e in range(2):
                    exp[k * 64 + j * 8 + phys * 2 + e] = base + k * 64 + j * 8 + chunk * 2 + e
    ok = bool((img == exp).all())
    print("SWIZZLE64 pattern matches assumption:", ok, flush=True)
    if not ok:
        print(" plane0 rows (got):")
        print(img[:64].reshape(8, 8) - base)
    return ok


def parity(ctx, dims=(32, 32, 32), boxes=None, seed=1, tag=""):
    dec = O.Decomp(dims, 8)
    nb = dec.nbricks
    rng = np.random.default_rng(seed)
    inp = rng.random(nb * 512)
    coeff = rng.random(129)
    d_in = ctx.upload(inp)
    d_out = ctx.malloc(inp.nbytes)
    plan = capi.Plan(ctx, dec.grid, dec.adj, boxes)
    if boxes is None:
        blist = dec.grid.reshape(-1)
    else:
        blist = np.concatenate([dec.grid[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]].reshape(-1) for lo, hi in boxes])
    allok = True
    for w in STENCILS:
        cf = coeff if w == 0 else None
        exp = np.full(nb * 512, -7.0)
        O.brick_sweep(w, dec.adj, blist, inp, exp, 512, 512, cf)
        res = {}
        for name, fn in (("naive", plan.stencil_naive), ("march", plan.stencil)):
            ctx.memset(d_out, 0, inp.nbytes)
            ctx.h2d(d_out, np.full(nb * 512, -7.0))
            ctx.sync()
            fn(w, d_in, d_out, 512, 512, cf)
            ctx.sync()
            got = ctx.download(d_out, nb * 512)
            res[name] = relerr(got, exp)
            if not res[name] <= 1e-12:
                bad = np.argwhere(np.abs(got - exp) > 1e-12 * np.abs(exp)).reshape(-1)
                bb = np.unique(bad // 512)
                print("   %s mismatch: %d elems in %d bricks; first elem %d (brick %d k%d j%d i%d) got %r exp %r"
                      % (name, bad.size, bb.size, bad[0], bad[0] // 512, (bad[0] % 512) // 64,
                         (bad[0] % 64) // 8, bad[0] % 8, got[bad[0]], exp[bad[0]]))
        ok = all(v <= 1e-12 for v in res.values())
        allok = allok and ok
        print("PARITY%s dims=%s stencil=%d %s %s" % (tag, dims, w, json.dumps(res), "OK" if ok else "FAIL"), flush=True)
    plan.close()
    ctx.free(d_in)
    ctx.free(d_out)
    return allok


def timing(ctx, n=512, reps=10):
    dec = O.Decomp((n, n, n), 8)
    nb = dec.nbricks
    rng = np.random.default_rng(3)
    inp = rng.random(nb * 512)
    d_a = ctx.upload(inp)
    d_b = ctx.malloc(inp.nbytes)
    ctx.memset(d_b, 0, inp.nbytes)
    plan = capi.Plan(ctx, dec.grid, dec.adj)
    coeff = rng.random(7)
    g = dec.grid.shape[0]
    for w in STENCILS:
        cf = coeff if w == 0 else None
        for _ in range(3):
            plan.stencil(w, d_a, d_b, 512, 512, cf)
            plan.stencil(w, d_b, d_a, 512, 512, cf)
        ctx.timer_start(0)
        for _ in range(reps):
            plan.stencil(w, d_a, d_b, 512, 512, cf)
            plan.stencil(w, d_b, d_a, 512, 512, cf)
        ctx.timer_stop(0)
        ms = ctx.timer_ms(0) / (2 * reps)
        pts = g ** 3 * 512
        print("TIMING n=%d stencil=%d P=%s ms/sweep=%.4f  Gpt/s(all bricks)=%.1f  GB/s(16B/pt)=%.0f  GStencil/s(interior)=%.1f"
              % (n, w, os.environ.get("BRICKS_PLANES", "4"), ms, pts / ms * 1e-6, pts * 16 / ms * 1e-6,
                 n ** 3 / ms * 1e-6), flush=True)
        # re-seed so values do not decay to denormals over many sweeps
        ctx.h2d(d_a, inp)
        ctx.sync()
    plan.close()
    ctx.free(d_a)
    ctx.free(d_b)


def main():
    quick = "--quick" in sys.argv
    t0 = time.time()
    ctx = capi.Context(0)
    out = {}
    if "--no-probes" not in sys.argv:
        probes(ctx, out)
        swizzle_check(ctx)
    ok = parity(ctx, (32, 32, 32))
    ok = parity(ctx, (16, 24, 40), tag="-ragged") and ok
    # odd-width boxes: interior box and two shell slabs
    ok = parity(ctx, (32, 32, 32), boxes=[((1, 1, 1), (4, 5, 5)), ((0, 0, 0), (6, 6, 1)), ((5, 0, 1), (6, 6, 6))],
                tag="-boxes") and ok
    print("PARITY ALL", "OK" if ok else "FAIL", "t=%.1fs" % (time.time() - t0), flush=True)
    if not quick:
        timing(ctx, 512)
    print("launches", capi.launch_count())
    ctx.close()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
