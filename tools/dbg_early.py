import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from bricklib_b200 import capi, weak
ctx = capi.Context(0)
for which in (25, 13, 7):
    sub = weak.WeakSubdomain(ctx, (32, 40, 48), which)
    f = sub.host_field()
    for nsteps in (1, 2, 3):
        res = []
        for early in (False, True):
            sub.upload(f)
            for _ in range(nsteps):
                sub.step(overlap=True, early=early)
            ctx.sync()
            res.append(sub.download())
        n = int(sub.dec.sep_pos[1]) * 512
        a, b = res[0][:n].reshape(-1, 512), res[1][:n].reshape(-1, 512)
        bad = np.nonzero((a != b).any(axis=1))[0]
        print(which, nsteps, "bad bricks", len(bad), bad[:10], "sep", sub.dec.sep_pos, "maxdiff", np.abs(a - b).max())
        if len(bad):
            g = sub.dec.grid
            pos = [tuple(int(x) for x in np.argwhere(g == bb)[0]) for bb in bad[:12]]
            print("  grid pos (k,j,i):", pos, "gdims", g.shape)
    sub.close()
