import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bricklib_b200 import capi
ctx = capi.Context(0)
nb = 4096
f = capi.uniform(nb * 512, seed=1)
d = ctx.upload(f)
for mode, name in ((0, "one lane, loop"), (1, "one lane, 8 unrolled"), (2, "12 lanes, divergent loop")):
    for nops in (24, 240):
        a, b = capi.probe_tma_issue(ctx, d, nb, mode, nops)
        print("TMAISSUE mode=%d (%s) nops=%d: issue %.1f cycles/op, until landed %.1f cycles/op" % (mode, name, nops, a, b))
