"""Worker of tests/test_4d.py (GPU part): runs the 4-D kernels of libbricks_b200.so through the C ABI and
compares them with the oracle (oracle/oracle4d.py), in a process of its own so that a device fault in these
kernels cannot take the CUDA context of the other GPU tests with it.  Prints one JSON object.

    python tests/gpu4d_worker.py parity | period | full
"""
import json
import os
import sys
import zlib

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)

from bricklib_b200 import capi, synth  # noqa: E402
from oracle import oracle4d as D  # noqa: E402

SENTINEL = -7.0


def relerr(got, exp):
    return float(np.max(np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)))


def run_sweep(ctx, plan, fn, inp, n, s):
    d_in = ctx.upload(inp)
    d_out = ctx.upload(np.full(n, SENTINEL))
    fn(d_in, d_out, s)
    ctx.sync()
    got = ctx.download(d_out, n)
    ctx.free(d_in)
    ctx.free(d_out)
    return got


def parity(ctx):
    """every sweep region of an exchange period, both kernels, three shapes (one with odd line counts),
    both numberings; then a grid whose cells were permuted, so that consecutive cells of a line are NOT
    neighbours and the marching kernel has to take its reload paths"""
    out = []
    for dims, pa in [((16, 16, 16, 16), False), ((16, 24, 16, 20), False), ((20, 20, 28, 16), True)]:
        dec = capi.Decomp4D(dims, 8, pa)
        n = dec.nbricks * 256
        inp = synth.uniform(n, seed=21)
        plan = capi.Plan4D(ctx, dec.grid, dec.adj)
        for s in (0, 1, 2, 3):
            exp = np.full(n, SENTINEL)
            D.brick_sweep(dec.adj, D.sweep_list(dec.grid, s), inp, exp)
            for name, fn in (("line", plan.sweep), ("naive", plan.sweep_naive)):
                got = run_sweep(ctx, plan, fn, inp, n, s)
                out.append({"dims": dims, "page_align": pa, "skip": s, "kernel": name, "relerr": relerr(got, exp),
                            "untouched_ok": bool(np.array_equal(got == SENTINEL, exp == SENTINEL)),
                            "cells": plan.ncells(s)})
        plan.close()
        dec.close()
    dec = capi.Decomp4D((16, 24, 16, 20), 8, False)
    rng = np.random.default_rng(5)
    grid = rng.permutation(dec.grid.reshape(-1)).reshape(dec.grid.shape)
    n = dec.nbricks * 256
    inp = synth.uniform(n, seed=22)
    plan = capi.Plan4D(ctx, grid, dec.adj)
    for s in (0, 2):
        exp = np.full(n, SENTINEL)
        D.brick_sweep(dec.adj, D.sweep_list(grid, s), inp, exp)
        for name, fn in (("line", plan.sweep), ("naive", plan.sweep_naive)):
            got = run_sweep(ctx, plan, fn, inp, n, s)
            out.append({"dims": "permuted grid", "skip": s, "kernel": name, "relerr": relerr(got, exp),
                        "untouched_ok": bool(np.array_equal(got == SENTINEL, exp == SENTINEL)), "cells": plan.ncells(s)})
    # a skip that leaves no brick: no launch, nothing written, no error
    for name, fn in (("line", plan.sweep), ("naive", plan.sweep_naive)):
        got = run_sweep(ctx, plan, fn, inp, n, 5)
        out.append({"dims": "empty box", "skip": 5, "kernel": name, "relerr": 0.0,
                    "untouched_ok": bool((got == SENTINEL).all()), "cells": plan.ncells(5)})
    plan.close()
    dec.close()
    return out


def period(ctx):
    """brick_func of weak/main-4d.cpp:247-277 on one periodic rank: exchange, then the ST_ITER = 8 sweeps
    with the shrinking regions; the oracle does the same with its own exchange and sweeps"""
    dims = (16, 24, 16, 20)
    dec = capi.Decomp4D(dims, 8, False)
    odec = D.DecompND(dims, page_align=False)
    n = dec.nbricks * 256
    a = synth.uniform(n, seed=23)
    b = np.zeros(n)
    dec.rank_map((1, 1, 1, 1), (0, 0, 0, 0))
    ex = capi.Exchange(ctx, 0, *dec.exchange_plan(256))
    plan = capi.Plan4D(ctx, dec.grid, dec.adj)
    d_a, d_b = ctx.upload(a), ctx.upload(b)
    oa, ob = a.copy(), b.copy()
    for _ in range(2):  # two periods: the second starts from exchanged, swept data
        ex.run(d_a, capi.STREAM_COMPUTE)
        D.exchange([oa], odec, (1, 1, 1, 1))
        src, dst, osrc, odst = d_a, d_b, oa, ob
        for skip in D.sweep_skips():
            plan.sweep(src, dst, skip // 4)
            D.brick_sweep(odec.adj, D.sweep_list(odec.grid, skip // 4), osrc, odst)
            src, dst, osrc, odst = dst, src, odst, osrc
    ctx.sync()
    got = ctx.download(d_a, n)
    interior = D.sweep_list(dec.grid, 2)
    g, e = got.reshape(-1, 256)[interior], oa.reshape(-1, 256)[interior]
    res = {"relerr_interior": relerr(g, e), "relerr_all": relerr(got, oa), "nregions": int(dec.nregions),
           "interior_bricks": int(len(interior))}
    for p in (d_a, d_b):
        ctx.free(p)
    ex.close()
    plan.close()
    return res


def full(ctx, dom=64, reps=20):
    """the largest case: a dom^4 subdomain (+ghost): every brick against the oracle, the two kernels against
    each other, and the time of one sweep"""
    dims = (dom,) * 4
    dec = capi.Decomp4D(dims, 8, False)
    n = dec.nbricks * 256
    inp = synth.uniform(n, seed=24)
    plan = capi.Plan4D(ctx, dec.grid, dec.adj)
    d_in = ctx.upload(inp)
    d_out = ctx.upload(np.zeros(n))
    d_nv = ctx.upload(np.zeros(n))
    plan.sweep(d_in, d_out, 0)
    plan.sweep_naive(d_in, d_nv, 0)
    ctx.sync()
    got, nv = ctx.download(d_out, n), ctx.download(d_nv, n)
    exp = np.zeros(n)
    D.brick_sweep(dec.adj, D.sweep_list(dec.grid, 0), inp, exp)
    res = {"dims": dims, "nbricks": int(dec.nbricks), "relerr": relerr(got, exp), "relerr_naive": relerr(nv, exp),
           "line_vs_naive": relerr(got, nv), "crc": "%08x" % zlib.crc32(got.tobytes())}
    cells = plan.ncells(0)
    for name, fn in (("line", plan.sweep), ("naive", plan.sweep_naive)):
        for _ in range(3):
            fn(d_in, d_out, 0)
        ctx.timer_start(0)
        for _ in range(reps):
            fn(d_in, d_out, 0)
        ctx.timer_stop(0)
        ms = ctx.timer_ms(0) / reps
        res[name + "_ms"] = ms
        res[name + "_GBps"] = cells * 256 * 16 / ms / 1e6
        res[name + "_GStencil"] = cells * 256 / ms / 1e6
    for p in (d_in, d_out, d_nv):
        ctx.free(p)
    plan.close()
    return res


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "parity"
    ctx = capi.Context(0)
    if what == "parity":
        res = parity(ctx)
    elif what == "period":
        res = period(ctx)
    else:
        res = full(ctx, int(sys.argv[2]) if len(sys.argv) > 2 else 64)
    ctx.close()
    print("RESULT " + json.dumps(res))


if __name__ == "__main__":
    main()
