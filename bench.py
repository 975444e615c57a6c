#!/usr/bin/env python
"""bench.py -- fp64 GStencil/s of the brick stencil hot path on N B200s (one rank per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload weak125|weak7|weak13|weak25]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's own CPU path (oracle/_ref)

One "step" is the hot loop of the reference's weak driver (weak/main.cu:246-287): one halo
exchange of the 512^3 subdomain each GPU owns, then ST_ITER stencil sweeps (4 for the 125-point
cube, 8 for the 7-point star).  value = interior points x sweeps x ranks / time, in GStencil/s
(weak/main.cu:315-317).  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {"weak125": 125, "weak7": 7, "weak13": 13, "weak25": 25}
SUBDOMAIN = (512, 512, 512)
BYTES_PER_POINT = 16.0  # one 8-byte read + one 8-byte write per point per sweep


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_reference_rate(which, budget_s=15.0, dims=(256, 256, 256)):
    """The reference's own CPU brick kernel (codegen, OpenMP, all host threads; oracle/_ref) on a
    bounded sample of the workload: full sweeps over a `dims` subdomain incl. ghost bricks."""
    from oracle import oracle as O
    from oracle import refload as R
    if not R.available():
        raise RuntimeError("oracle/_ref is not built for this host")
    dec = O.Decomp(dims, 8)
    n = dec.nbricks * 512
    a, b = R.aligned_empty(n), R.aligned_empty(n)
    from bricklib_b200 import synth
    a[:] = synth.uniform(n, seed=1)
    b[:] = 0.0
    blist = dec.grid.reshape(-1)
    t1 = R.brick_sweep_timed(which, dec.adj, blist, a, b, 512, None, 1)
    reps = int(max(2, min(400, budget_s / max(t1, 1e-4))))
    reps += reps & 1
    t = R.brick_sweep_timed(which, dec.adj, blist, a, b, 512, None, reps)
    pts = float(np.prod(dims))
    return {"value": pts / t * 1e-9, "unit": "GStencil/s", "cores": R.max_threads(), "kind": "reference",
            "sample": "%d sweeps of the %d-point stencil over one %dx%dx%d subdomain (+ghost bricks), "
                      "reference codegen brick kernel (%s), %.1f ms/sweep" % (reps, which if which else 7, dims[0], dims[1],
                                                                               dims[2], R.vsvec(), t * 1e3)}
