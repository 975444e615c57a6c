#!/bin/bash
# one GPU: fp32 bricks -- parity tests, fp64 regression by CRC (the product kernels must read d2325f1e /
# 1ad50a39 / e4af5d65 / 5fbf54b1 for 125 / 25 / 13 / 7), and the float kernels' sweep times
cd "$(dirname "$0")/.."
O=gpurun_out/fp32; mkdir -p $O
timeout 200 python -m pytest tests/test_fp32.py -x -q -m gpu 2>&1 | tail -12 > $O/pytest_fp32.log; cat $O/pytest_fp32.log
VARIANT=product timeout 60 python tools/variant_check.py 125,25,13,7 2>&1 | grep VCHECK | tee $O/fp64_crc.log
timeout 90 python - <<'PY' 2>&1 | tee $O/fp32_timing.log
import numpy as np, sys
sys.path.insert(0, ".")
from bricklib_b200 import capi
ctx = capi.Context(0)
dec = capi.Decomp((512, 512, 512), 8)
nb = dec.nbricks
x = capi.uniform(nb * 512, seed=3).astype(np.float32)
d_a, d_b = ctx.upload(x), ctx.upload(np.zeros(nb * 512, np.float32))
plan = capi.Plan(ctx, dec.grid, dec.adj)
for w in (7, 13, 25, 125):
    for _ in range(2):
        plan.stencil_f32(w, d_a, d_b); plan.stencil_f32(w, d_b, d_a)
    ctx.timer_start(0)
    for _ in range(10):
        plan.stencil_f32(w, d_a, d_b); plan.stencil_f32(w, d_b, d_a)
    ctx.timer_stop(0)
    ms = ctx.timer_ms(0) / 20
    print("F32TIME stencil=%d ms=%.4f GB/s(8 B/pt, all bricks)=%.0f interior GStencil/s=%.1f" % (w, ms, 66**3 * 512 * 8 / ms * 1e-6, 512**3 / ms * 1e-6))
PY
