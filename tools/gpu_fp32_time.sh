#!/bin/bash
# one GPU: sweep times of the float kernels for the product library and for experiment builds
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/fp32
for lib in libbricks_b200.so "$@"; do
  echo "== $lib"
  BRICKS_LIB=$PWD/bricklib_b200/lib/$lib timeout 60 python - <<'PY' 2>&1 | grep F32TIME
import numpy as np, sys, zlib
sys.path.insert(0, ".")
from bricklib_b200 import capi
ctx = capi.Context(0)
dec = capi.Decomp((512, 512, 512), 8)
nb = dec.nbricks
x = capi.uniform(nb * 512, seed=3).astype(np.float32)
d_a, d_b = ctx.upload(x), ctx.upload(np.zeros(nb * 512, np.float32))
plan = capi.Plan(ctx, dec.grid, dec.adj)
for w in (7, 13, 25, 125):
    ctx.h2d(d_a, x); ctx.sync()
    plan.stencil_f32(w, d_a, d_b); ctx.sync()
    crc = zlib.crc32(ctx.download(d_b, nb * 512, np.float32).tobytes())
    for _ in range(2):
        plan.stencil_f32(w, d_a, d_b); plan.stencil_f32(w, d_b, d_a)
    ctx.timer_start(0)
    for _ in range(10):
        plan.stencil_f32(w, d_a, d_b); plan.stencil_f32(w, d_b, d_a)
    ctx.timer_stop(0)
    ms = ctx.timer_ms(0) / 20
    print("F32TIME stencil=%d crc=%08x ms=%.4f GB/s(8 B/pt, all bricks)=%.0f interior GStencil/s=%.1f" % (w, crc, ms, 66**3 * 512 * 8 / ms * 1e-6, 512**3 / ms * 1e-6))
PY
done | tee gpurun_out/fp32/fp32_caps.log
