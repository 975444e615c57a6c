#!/usr/bin/env python
"""bench.py -- fp64 GStencil/s of the brick stencil hot path on N B200s (one rank per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload weak125|weak7|weak13|weak25|strong125|strong7]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's own CPU path (oracle/_ref)

One "step" is the hot loop of the reference's weak driver (weak/main.cu:246-287): one halo
exchange of the 512^3 subdomain each GPU owns, then ST_ITER stencil sweeps (4 for the 125-point
cube, 8 for the 7-point star).  value = interior points x sweeps x ranks / time, in GStencil/s
(weak/main.cu:315-317).  Prints ONE JSON line (rank 0).

What the line carries beyond the contract keys:
  parity     before anything is timed, at EVERY N: one bare halo exchange and two weak steps (7- and
             125-point) on a small subdomain per rank through the same WeakSubdomain code the timed
             loop uses, gathered to rank 0 and compared with the oracle's BrickDecomp::exchange
             restatement (bit-exact) and with the global periodic array computation (<= 1e-12).
             The oracle is used here as the checker only.  A failure exits non-zero.
  roofline   the dominant kernel (one sweep over the 66^3-brick grid): `frac` counts the bytes the
             launch really moves (ghost bricks included, = DRAM utilisation, cross-checked by ncu
             dram bytes), `frac_interior` counts only the 512^3 interior points SURVEY.md 8(d)
             calls algorithmic (ghost-brick recompute is overhead).
  kernels    the same two fractions for every stencil of the path (7/13/25/125-point).
  strong     BASELINE.json configs[4] at this N: 1024^3 in 128^3 subdomains, 125-point, a few steps.
  cpu_baseline  N = 1 only: the reference's codegen brick kernel (oracle/_ref) on all host cores,
             full 512^3 sweeps.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {"weak125": 125, "weak7": 7, "weak13": 13, "weak25": 25, "strong125": 125, "strong7": 7}
SUBDOMAIN = (512, 512, 512)
STRONG_DOM, STRONG_SDOM = 1024, 128  # BASELINE.json configs[4]: 1024^3 global grid in 128^3 subdomains
BYTES_PER_POINT = 16.0  # one 8-byte read + one 8-byte write per point per sweep
ST_ITER = {125: 4, 7: 8, 13: 4, 25: 2}  # stencils/fake.h:39-344 (GZ / radius)
REL_TOL = 1e-12  # BASELINE.json north_star: stencil output vs the reference CPU path
PARITY_DIMS = (64, 64, 64)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_source_hash():
    """sha1 over the CUDA sources of the sweep kernels: ties profiles/traffic.json (ncu dram bytes)
    to the kernel version that is actually benchmarked."""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "bricklib_b200", "csrc")
    for name in ("march_dev.cuh", "march_steps.cuh", "stencil_march.cu"):
        with open(os.path.join(d, name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def ncu_traffic(which):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full
    capture of THIS kernel version, else None (a capture of another version is not reported)."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(tpath) as f:
            t = json.load(f)
    except Exception:  # noqa: BLE001
        return None, "no profiles/traffic.json"
    if t.get("kernel_source_hash") != kernel_source_hash():
        return None, "profiles/traffic.json was captured from another kernel version (%s)" % t.get("kernel_source_hash")
    return t.get(str(which)), t.get("source")


def dims_create(nnodes, ndims=3):
    """MPI_Dims_create for the process counts of one box (weak/args.cpp:105-130): 8 -> 2x2x2,
    4 -> 2x2x1, 2 -> 2x1x1.  Pure Python so that the reference arm needs nothing of the product."""
    dims = [1] * ndims
    n, f = nnodes, 2
    factors = []
    while n > 1:
        while n % f == 0:
            factors.append(f)
            n //= f
        f += 1
    for p in sorted(factors, reverse=True):
        dims[dims.index(min(dims))] *= p
    return sorted(dims, reverse=True)


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons during the timed region (B200_PROFILING.md), through NVML in
    this process.  Forking nvidia-smi every few milliseconds instead stalls the driver calls of the
    rank that does it (fork of a process with a CUDA context) and, on a periodic process grid, every
    rank that waits for it; nvidia-smi stays as the fallback where NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = [0x8, 0x40, 0x20, 0x4]  # nvmlClocksEventReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, index, pci_bus_id=None):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.nvml = self.handle = None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if pci_bus_id:
                try:
                    h = pynvml.nvmlDeviceGetHandleByPciBusId(pci_bus_id.encode() if isinstance(pci_bus_id, str) else pci_bus_id)
                except Exception:  # noqa: BLE001
                    h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.nvml, self.handle = pynvml, h
        except Exception:  # noqa: BLE001
            self.nvml = None

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        mhz = float(n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM))
        try:
            mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(h))
        except Exception:  # noqa: BLE001 -- older binding name
            mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(h))
        return [mhz, self.max_mhz] + [bool(mask & b) for b in self.BITS]

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        parts = [x.strip() for x in out.strip().split(",")]
        if len(parts) < 6:
            return None
        return [float(parts[0]), float(parts[1])] + [p.lower().startswith("active") for p in parts[2:6]]

    def run(self):
        while not self.stop_flag:
            try:
                row = self._sample_nvml() if self.nvml else self._sample_smi()
                if row:
                    self.rows.append(row)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.004 if self.nvml else 0.05)

    def reset(self):
        """forget what was sampled so far (the sampler is started before the timed region)"""
        self.rows = []

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for i, n in enumerate(self.NAMES) if any(r[2 + i] for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.rows[0][1], "reasons": reasons,
                "samples": len(self.rows), "source": "nvml" if self.nvml else "nvidia-smi"}


# ------------------------------------------------------------------------------------------
# CPU legs: the reference's own code (oracle/_ref).  Nothing of bricklib_b200's library is loaded.

def use_all_host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the CPU legs must use the
    host's cores whatever the launcher exported.  Sets the OpenMP thread count of THIS thread's
    parallel regions to the size of the process's affinity mask (libgomp is the runtime
    oracle/_ref links) and returns it."""
    import ctypes
    n = len(os.sched_getaffinity(0))
    os.environ["OMP_NUM_THREADS"] = str(n)  # for a runtime that has not started yet
    ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL).omp_set_num_threads(n)
    return n


def synth_uniform(n, seed):
    """bricklib_b200/synth.py is pure numpy (the counter-based U[0,1) generator of
    include/bricks/multiarray.h); importing it does not load the CUDA library."""
    from bricklib_b200 import synth
    return synth.uniform(n, seed=seed)


def cpu_reference_rate(which, budget_s=15.0, dims=SUBDOMAIN):
    """The reference's own CPU brick kernel (codegen, OpenMP, all host threads; oracle/_ref) on a
    bounded sample of the workload: full sweeps over a `dims` subdomain incl. ghost bricks."""
    from oracle import oracle as O
    from oracle import refload as R
    if not R.available():
        raise RuntimeError("oracle/_ref is not built for this host")
    threads = use_all_host_threads()
    if R.max_threads() != threads:
        raise RuntimeError("OpenMP runs %d threads, the affinity mask has %d" % (R.max_threads(), threads))
    dec = O.Decomp(dims, 8)
    n = dec.nbricks * 512
    a, b = R.aligned_empty(n), R.aligned_empty(n)
    a[:] = synth_uniform(n, 1)
    b[:] = 0.0
    blist = dec.grid.reshape(-1)
    t1 = R.brick_sweep_timed(which, dec.adj, blist, a, b, 512, None, 2)
    reps = int(max(2, min(400, budget_s / max(t1, 1e-4))))
    reps += reps & 1
    t = R.brick_sweep_timed(which, dec.adj, blist, a, b, 512, None, reps)
    pts = float(np.prod(dims))
    return {"value": pts / t * 1e-9, "unit": "GStencil/s", "cores": threads, "kind": "reference",
            "subdomain": list(dims),
            "sample": "%d sweeps of the %d-point stencil over one %dx%dx%d subdomain (+ghost bricks), "
                      "reference codegen brick kernel (%s) under OpenMP with %d threads, %.1f ms/sweep"
                      % (reps, which if which else 7, dims[0], dims[1], dims[2], R.vsvec(), threads, t * 1e3)}


def run_reference(args, which):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref: its
    codegen brick kernel under OpenMP, all host threads) on this arm's metric/config.  A step is
    what one rank of the workload does: one periodic self exchange + ST_ITER sweeps over the
    512^3 subdomain (strong: over one 128^3 subdomain).  Rank 0 only; maps no product library."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as O
    from oracle import refload as R
    threads = use_all_host_threads()
    st_iter = ST_ITER[which]
    strong = args.workload.startswith("strong")
    # strong: a step of the sample = one subdomain of the decomposition (its 26 periodic neighbours
    # are copies of itself, as in strong/main.cu on one rank with one subdomain)
    dims = (args.sdom,) * 3 if strong else SUBDOMAIN
    dec = O.Decomp(dims, 8)
    n = dec.nbricks * 512
    a, b = R.aligned_empty(n), R.aligned_empty(n)
    a[:] = synth_uniform(n, 1)
    b[:] = 0.0
    blist = np.ascontiguousarray(dec.grid.reshape(-1), np.uint32)
    adjf = np.ascontiguousarray(dec.adj.reshape(-1), np.uint32)
    sweep = getattr(R.lib(), "ref%d_brick_sweep" % which)
    # a step = one (self) exchange + ST_ITER sweeps, as weak/main.cpp:172-213 on one rank
    regions = [((int(g["pos"]) + int(g["first_pad"])) * 512, (int(k["pos"]) + int(k["first_pad"])) * 512,
                (int(g["len"]) - int(g["first_pad"]) - int(g["last_pad"])) * 512) for g, k in zip(dec.ghost, dec.skin)]

    def step():
        for d0, s0, cnt in regions:  # BrickDecomp::exchange on one periodic rank: ghost[i] <- skin[i]
            a[d0:d0 + cnt] = a[s0:s0 + cnt]
        src, dst = a, b
        for _ in range(st_iter):  # native fold order of this build: layout does not matter for timing
            sweep(adjf, dec.nbricks, blist, blist.size, src, dst, 512, None)
            src, dst = dst, src
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = float(np.prod(dims)) * st_iter / dt * 1e-9
    sample = ("%d steps x (periodic self exchange + %d sweeps) over one %d^3 subdomain (+ghost bricks): what one rank "
              "of the workload does per step; codegen brick kernel %s under OpenMP with %d threads (affinity mask %d, "
              "launcher exported OMP_NUM_THREADS=%s)"
              % (args.steps, st_iter, dims[0], R.vsvec(), R.max_threads(), threads, args.launcher_omp))
    line = {"impl": "reference", "metric": "fp64 GStencil/s", "value": val, "unit": "GStencil/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": make_config(which, args.gpus, (args.dom, args.sdom) if strong else None),
            "cpu_baseline": {"value": val, "unit": "GStencil/s", "cores": R.max_threads(), "kind": "reference",
                             "subdomain": list(dims), "sample": sample},
            "e2e": {"value": val, "unit": "GStencil/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def make_config(which, ngpus, strong=None):
    if strong:
        dom, sdom = strong
        g = sdom // 8 + 2
        return {"workload": ("strong-scaling step: %d-point %s stencil, fp64, %d^3 global periodic grid in %d^3 subdomains "
                             "(Z-Morton ids, contiguous ranges per GPU), 8x8x8 bricks, 1 halo exchange + %d sweeps per step"
                             % (which, "cube" if which == 125 else "star", dom, sdom, ST_ITER[which])),
                "global_grid": [dom] * 3, "subdomain": [sdom] * 3, "subdomains": (dom // sdom) ** 3, "brick": [8, 8, 8],
                "ghost_depth": 8, "st_iter": ST_ITER[which],
                "l2_policy": "working set %.2f GB per sweep over all GPUs >> 126 MB L2 (inputs larger than L2)"
                             % ((dom // sdom) ** 3 * g ** 3 * 4096 * 2 / 1e9)}
    return {"workload": workload_name(which), "subdomain_per_gpu": list(SUBDOMAIN), "brick": [8, 8, 8],
            "ghost_depth": 8, "st_iter": ST_ITER[which], "process_grid": dims_create(ngpus),
            "l2_policy": "working set 2.36 GB per GPU per sweep >> 126 MB L2 (inputs larger than L2)"}


def workload_name(which):
    return ("weak-scaling step: %d-point %s stencil, fp64, 512^3 subdomain per GPU in 8x8x8 bricks "
            "(BrickDecomp layout), 1 periodic halo exchange + %d sweeps per step"
            % (which, "cube" if which == 125 else "star", ST_ITER[which]))


# ------------------------------------------------------------------------------------------
# GPU arm

def kernel_probe(ctx, sub, which, reps=10):
    """Average duration of the dominant kernel (one full-grid sweep, 66^3 bricks), CUDA events on
    the stream it is launched on, back to back ping-pong launches."""
    for _ in range(2):
        sub.sweep_full(which, sub.d_a, sub.d_b)
        sub.sweep_full(which, sub.d_b, sub.d_a)
    ctx.timer_start(1)
    for _ in range(reps):
        sub.sweep_full(which, sub.d_a, sub.d_b)
        sub.sweep_full(which, sub.d_b, sub.d_a)
    ctx.timer_stop(1)
    return ctx.timer_ms(1) / (2 * reps)


def kernel_fractions(ms, gdims, dims, peak):
    """The two ways of charging a sweep against the HBM roof (module docstring)."""
    launch_pts = float(gdims[0] * gdims[1] * gdims[2] * 512)
    interior = float(np.prod(dims))
    achieved = launch_pts * BYTES_PER_POINT / (ms * 1e-3) * 1e-9
    achieved_int = interior * BYTES_PER_POINT / (ms * 1e-3) * 1e-9
    return {"achieved": achieved, "frac": achieved / peak, "achieved_interior": achieved_int,
            "frac_interior": achieved_int / peak, "frac_of_8TBps_nominal": achieved / 8000.0,
            "frac_interior_of_8TBps_nominal": achieved_int / 8000.0,
            "interior_gstencil_per_s_kernel_only": interior / (ms * 1e-3) * 1e-9,
            "algorithmic_bytes_per_launch": launch_pts * BYTES_PER_POINT,
            "algorithmic_bytes_interior": interior * BYTES_PER_POINT}


def init_ranks(args):
    """one process per GPU: context on LOCAL_RANK, NCCL communicator shared through torch.distributed"""
    import torch
    import torch.distributed as dist
    from bricklib_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d"
                             % (args.gpus, args.gpus))
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    ctx = capi.Context(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.frombuffer(bytearray(capi.unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))

    def barrier():
        ctx.sync()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather(arr):
        """every rank's numpy array (same shape on all ranks) on every rank"""
        if world == 1:
            return [arr]
        t = torch.from_numpy(np.ascontiguousarray(arr)).cuda()
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t)
        return [o.cpu().numpy() for o in outs]

    def allgather_bytes(b):
        """every rank's byte string (same length on all ranks), in rank order"""
        if world == 1:
            return [b]
        t = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t)
        return [bytes(o.cpu().numpy().tobytes()) for o in outs]

    gather.bytes = allgather_bytes
    return ctx, rank, world, local, dist, barrier, max_over_ranks, gather


def parity_block(ctx, rank, world, gather, overlap, early, kind, unsplit=False, p2p=False):
    """Correctness evidence inside the run the driver times (module docstring).  Same code path as
    the timed loop -- WeakSubdomain over the C ABI, NCCL between the ranks -- on PARITY_DIMS per
    rank.  Checked on rank 0 against (1) the oracle's restatement of BrickDecomp::exchange
    (include/brick-mpi.h:466-495) on `world` virtual ranks, bit for bit, and (2) the stencil applied
    to the assembled GLOBAL periodic array (what weak/main.cu:321-323 checks per rank)."""
    from bricklib_b200 import capi, weak
    dims = PARITY_DIMS
    procdims = capi.cart_dims(world)
    out = {"nranks": world, "subdomain_per_rank": list(dims), "process_grid": list(procdims), "tolerance": REL_TOL,
           "schedule": {"overlap": overlap, "early_exchange": early, "exchange": kind, "unsplit": unsplit, "p2p": p2p},
           "checker": "oracle/oracle.py exchange() (bit-exact) and the global periodic array through the oracle's "
                      "array stencil"}
    ok = True
    O = odec = None
    if rank == 0:
        from oracle import oracle as O  # the checker; never on the measured path
        odec = O.Decomp(dims, 8)

    # (1) one bare exchange of a random storage
    sub = weak.WeakSubdomain(ctx, dims, 7, rank=rank, procdims=procdims, exchange=kind)
    if p2p:
        sub.enable_p2p(gather.bytes)
    nb = sub.dec.nbricks
    f = capi.uniform(nb * 512, seed=9000 + rank)
    sub.upload(f)
    sub.exchange_only()
    ctx.sync()
    got = gather(sub.download())
    fields = gather(f)
    gather(np.zeros(1))  # every rank is done with its peers' memory before anybody unmaps it
    sub.close()
    if rank == 0:
        exp = O.exchange([x.copy() for x in fields], odec, procdims)
        bit = all(np.array_equal(g.view(np.uint64), e.view(np.uint64)) for g, e in zip(got, exp))
        changed = int(sum(np.count_nonzero(g.view(np.uint64) != x.view(np.uint64)) for g, x in zip(got, fields)))
        out["exchange_bitexact"] = bool(bit)
        out["exchange_words_moved"] = changed
        ok = ok and bit and changed > 0

    # (2) two steps of the 7- and the 125-point stencil == periodic sweeps of the global array
    t = [x // 8 + 2 for x in dims]
    ext = [t[0] * 8, t[1] * 8, t[2] * 8]
    pk, pj, pi = procdims

    def where(r):  # physical block of rank r: populate() puts the "+d" neighbour at coordinate - 1
        c = capi.cart_coords(r, procdims)
        return [(-c[a]) % procdims[a] for a in range(3)]

    errs = {}
    nsteps = 2
    for which in (7, 125):
        sub = weak.WeakSubdomain(ctx, dims, which, rank=rank, procdims=procdims, exchange=kind)
        if p2p:
            sub.enable_p2p(gather.bytes)
        f0 = sub.host_field()
        sub.upload(f0)
        for _ in range(nsteps):
            sub.step(overlap=overlap, early=early, unsplit=unsplit)
        ctx.sync()
        res = gather(sub.download())
        f0s = gather(f0)
        gather(np.zeros(1))
        sub.close()
        if rank == 0:
            def interior_of(field):
                a = np.zeros((ext[2], ext[1], ext[0]))
                O.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], a, odec.grid, field)
                return a[8:-8, 8:-8, 8:-8]
            d2, d1, d0 = dims[2], dims[1], dims[0]
            glob = np.zeros((d2 * pk, d1 * pj, d0 * pi))
            for r in range(world):
                ck, cj, ci = where(r)
                glob[d2 * ck:d2 * (ck + 1), d1 * cj:d1 * (cj + 1), d0 * ci:d0 * (ci + 1)] = interior_of(f0s[r])
            cur = glob
            for _ in range(nsteps * ST_ITER[which]):
                pad = np.pad(cur, 8, mode="wrap")
                o = np.zeros_like(pad)
                O.array_sweep(which, pad, o, (8, 8 + cur.shape[0], 8, 8 + cur.shape[1], 8, 8 + cur.shape[2]))
                cur = o[8:-8, 8:-8, 8:-8].copy()
            worst = 0.0
            for r in range(world):
                ck, cj, ci = where(r)
                e = cur[d2 * ck:d2 * (ck + 1), d1 * cj:d1 * (cj + 1), d0 * ci:d0 * (ci + 1)]
                g = interior_of(res[r])
                worst = max(worst, float(np.max(np.abs(g - e) / np.maximum(np.abs(e), 1e-300))))
            errs[str(which)] = worst
            ok = ok and worst <= REL_TOL
    if rank == 0:
        out["step_max_rel_err"] = errs
        out["steps_checked"] = nsteps
        out["ok"] = bool(ok)
    return out


def run_gpu(args, which):
    from bricklib_b200 import capi, weak
    ctx, rank, world, local, dist, barrier, max_over_ranks, gather = init_ranks(args)

    procdims = capi.cart_dims(world)
    st_iter = capi.ST_ITER[which]
    # start the next step's exchange behind the last sweep's skin bricks: pays off when the
    # exchange crosses NVLink (N > 1); on one GPU the plain schedule is faster (measured)
    overlap = (not args.no_overlap) and (world > 1 or args.overlap)
    p2p = bool(args.p2p) and world > 1 and args.exchange != "packed"
    early = world > 1 and overlap and not args.no_early and (args.early or not p2p)
    parity = None
    if not args.no_parity:
        parity = parity_block(ctx, rank, world, gather, overlap, early, args.exchange, args.unsplit, p2p)
        bad = 0.0 if (rank != 0 or parity["ok"]) else 1.0
        if max_over_ranks(bad) > 0:
            if rank == 0:
                print(json.dumps({"error": "parity check failed", "parity": parity}), flush=True)
            return 3

    sub = weak.WeakSubdomain(ctx, SUBDOMAIN, which, rank=rank, procdims=procdims, exchange=args.exchange)
    p2p_note = None
    if p2p and not sub.enable_p2p(gather.bytes):
        p2p, p2p_note = False, "fell back to NCCL: %s" % sub.p2p_error
        early = world > 1 and overlap and not args.no_early  # NCCL's exchange wants the early schedule
    field = sub.host_field()
    sub.upload(field)

    # ---- device-resident timing: W warm-up steps, then exactly K steps
    for _ in range(max(args.warmup, 3)):
        sub.step(overlap=overlap, early=early, unsplit=args.unsplit)
    if args.trace:  # one traced step per rank, outside the timed region
        barrier()
        sub.step_obj.trace(True)
        sub.step(overlap=overlap, early=early, unsplit=args.unsplit)
        tl = sub.step_obj.trace_read()
        sub.step_obj.trace(False)
        with open("%s.rank%d.json" % (args.trace, rank), "w") as f:
            json.dump({"rank": rank, "nranks": world, "workload": args.workload, "exchange": sub.exchange_kind,
                       "early": early, "unsplit": args.unsplit,
                       "timeline_ms": [{"op": n, "start": a, "end": b} for n, a, b in tl]}, f, indent=1)
    sampler = ClockSampler(local) if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS") else None  # (diagnosis switch)
    if sampler:
        sampler.start()
    barrier()
    if sampler:
        sampler.reset()
    n0 = capi.launch_count()
    ctx.timer_start(0)
    for _ in range(args.steps):
        sub.step(overlap=overlap, early=early, unsplit=args.unsplit)
    ctx.stream_wait(capi.STREAM_COMPUTE, capi.STREAM_COMM)  # nothing of the K steps is left in flight
    ctx.timer_stop(0)
    ms = ctx.timer_ms(0)
    barrier()
    launches = capi.launch_count() - n0
    ms = max_over_ranks(ms)
    ms_step = ms / args.steps
    pts_step = float(np.prod(SUBDOMAIN)) * world * st_iter
    value = pts_step / (ms_step * 1e-3) * 1e-9

    # ---- dominant kernel: one full-grid sweep
    sub.upload(field)
    k_ms = kernel_probe(ctx, sub, which)
    peak, peak_src = measured_peak()
    traffic, traffic_src = ncu_traffic(which)
    roofline = {"bound": "hbm", "peak": peak, "unit": "GB/s", "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "kernel": "march_kernel (one sweep over 66^3 bricks)", "kernel_ms": k_ms,
                "kernel_source_hash": kernel_source_hash(),
                "definition": "frac = 16 B x every point the launch updates (66^3 bricks, ghost bricks included) / "
                              "kernel_ms / peak = DRAM utilisation; frac_interior = 16 B x 512^3 interior points / "
                              "kernel_ms / peak (SURVEY.md 8d: ghost-brick recompute is overhead, not counted)"}
    roofline.update(kernel_fractions(k_ms, sub.gdims, SUBDOMAIN, peak))

    # ---- end to end through the C ABI with HOST buffers: H2D of the step's input field from
    # pinned memory, exchange + ST_ITER sweeps, D2H of the step's result
    e2e = end_to_end(ctx, sub, field, args, barrier, max_over_ranks, pts_step)
    kernels = None
    if not args.no_all_stencils:
        kernels = all_stencils(ctx, sub, peak)
    # the sampler ran through every timed region (device-resident steps, kernel probes, e2e steps)
    clocks = sampler.summary() if sampler else None
    exchange_bytes = sub.exchange_bytes
    exchange_kind = sub.exchange_kind
    barrier()  # nobody unmaps a storage its peers may still be copying into
    sub.close()

    strong = None
    if not args.no_strong:
        try:
            strong = strong_block(ctx, rank, world, barrier, max_over_ranks, args, peak)
        except Exception as e:  # noqa: BLE001 -- the extra block must not take the headline line down
            strong = {"error": repr(e)}

    d4 = None
    if world == 1 and not args.no_4d:
        d4 = d4pt9_block(peak)

    line = None
    if rank == 0:
        line = {"metric": "fp64 GStencil/s", "value": value, "unit": "GStencil/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": make_config(which, world), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roofline, "parity": parity,
                "overlap": overlap, "early_exchange": early, "early_unsplit": bool(args.unsplit and early), "peer_memory_exchange": p2p, "peer_memory_exchange_note": p2p_note,
                "nccl_max_ctas": int(os.environ.get("BRICKS_NCCL_MAX_CTAS", "0")),
                "exchange_bytes_per_step_per_gpu": exchange_bytes, "exchange": exchange_kind}
        if kernels is not None:
            line["kernels"] = kernels
        if strong is not None:
            line["strong"] = strong
        if d4 is not None:
            line["d4pt9"] = d4
        if world == 1 and not args.no_cpu:
            try:
                line["cpu_baseline"] = cpu_reference_rate(which, budget_s=args.cpu_seconds)
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": "GStencil/s", "cores": 0, "kind": "reference",
                                        "sample": "unavailable: %r" % (e,)}
        elif world > 1:
            line["cpu_baseline"] = None  # timed at N = 1 only (the other ranks' host threads would compete)
    if world > 1:
        dist.barrier()
    if line:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def strong_block(ctx, rank, world, barrier, max_over_ranks, args, peak, which=125, steps=5):
    """BASELINE.json configs[4] next to the headline: strong scaling of the 125-point stencil,
    1024^3 in 128^3 subdomains shared out over the ranks, device-resident, `steps` steps."""
    from bricklib_b200 import capi, strong
    dom, sdom = args.dom, args.sdom
    sd = strong.StrongDomain(ctx, dom, sdom, which, rank=rank, nranks=world)
    sd.upload()
    for _ in range(3):
        sd.step(overlap=not args.no_overlap)
    barrier()
    ctx.timer_start(3)
    for _ in range(steps):
        sd.step(overlap=not args.no_overlap)
    ctx.timer_stop(3)
    ms = ctx.timer_ms(3)
    barrier()
    ms_step = max_over_ranks(ms) / steps
    for _ in range(2):
        sd.sweep_full(which, sd.in_a, sd.in_b)
        sd.sweep_full(which, sd.in_b, sd.in_a)
    ctx.timer_start(3)
    for _ in range(3):
        sd.sweep_full(which, sd.in_a, sd.in_b)
        sd.sweep_full(which, sd.in_b, sd.in_a)
    ctx.timer_stop(3)
    k_ms = ctx.timer_ms(3) / 6
    barrier()
    g = sd.dec.gdims
    fr = kernel_fractions(k_ms, [g[0] * sd.nsub, g[1], g[2]], (sdom * sd.nsub, sdom, sdom), peak)
    out = {"workload": make_config(which, world, (dom, sdom))["workload"],
           "value": float(dom) ** 3 * capi.ST_ITER[which] / (ms_step * 1e-3) * 1e-9, "unit": "GStencil/s",
           "ms_per_step": ms_step, "steps": steps, "warmup": 3, "scaling": "strong", "subdomains_rank0": sd.nsub,
           "exchange_bytes_sent_rank0": sd.exchange_bytes, "kernel_ms": k_ms, "frac": fr["frac"],
           "frac_interior": fr["frac_interior"], "ghost_overhead": float(g[0] * g[1] * g[2] * 512) / float(sdom) ** 3}
    sd.close()
    return out


def run_gpu_strong(args, which):
    """Strong scaling (strong/main.cu:397-540): the global dom^3 grid is fixed, its (dom/sdom)^3
    subdomains are shared out over the ranks.  value = dom^3 x ST_ITER / time of a step."""
    import ctypes as C
    from bricklib_b200 import capi, strong
    ctx, rank, world, local, dist, barrier, max_over_ranks, gather = init_ranks(args)
    dom, sdom = args.dom, args.sdom
    st_iter = capi.ST_ITER[which]
    sd = strong.StrongDomain(ctx, dom, sdom, which, rank=rank, nranks=world)
    # pinned host images of all local subdomain storages (input, result)
    nbytes = sd.nsub * sd.sub_bytes
    n = nbytes // 8
    ptrs = []
    for _ in range(2):
        p = C.c_void_p()
        capi.check(capi.lib().bricks_host_alloc(C.byref(p), nbytes))
        ptrs.append(p)
    h_in = np.ctypeslib.as_array((C.c_double * n).from_address(ptrs[0].value))
    h_out = np.ctypeslib.as_array((C.c_double * n).from_address(ptrs[1].value))
    per = sd.sub_bytes // 8
    for s_ in range(sd.nsub):
        h_in[s_ * per:(s_ + 1) * per] = sd.host_field(s_)
    ctx.h2d(sd.d_a, h_in)
    ctx.sync()
    for _ in range(max(args.warmup, 3)):
        sd.step(overlap=not args.no_overlap)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    barrier()
    if sampler:
        sampler.reset()
    n0 = capi.launch_count()
    ctx.timer_start(0)
    for _ in range(args.steps):
        sd.step(overlap=not args.no_overlap)
    ctx.timer_stop(0)
    ms = ctx.timer_ms(0)
    barrier()
    launches = capi.launch_count() - n0
    clocks = sampler.summary() if sampler else None
    ms_step = max_over_ranks(ms) / args.steps
    pts_step = float(dom) ** 3 * st_iter
    value = pts_step / (ms_step * 1e-3) * 1e-9

    # ---- dominant kernel: one sweep over every brick (ghost included) of all local subdomains
    ctx.h2d(sd.d_a, h_in)
    ctx.sync()
    for _ in range(2):
        sd.sweep_full(which, sd.in_a, sd.in_b)
        sd.sweep_full(which, sd.in_b, sd.in_a)
    reps = 5
    ctx.timer_start(1)
    for _ in range(reps):
        sd.sweep_full(which, sd.in_a, sd.in_b)
        sd.sweep_full(which, sd.in_b, sd.in_a)
    ctx.timer_stop(1)
    k_ms = ctx.timer_ms(1) / (2 * reps)
    g = sd.dec.gdims
    peak, peak_src = measured_peak()
    roofline = {"bound": "hbm", "peak": peak, "unit": "GB/s", "traffic": None, "peak_source": peak_src,
                "kernel": "march_kernel (one sweep over %d subdomains x %d^3 bricks, one launch)" % (sd.nsub, g[0]),
                "kernel_ms": k_ms, "ghost_overhead": float(g[0] * g[1] * g[2] * 512) / float(sdom) ** 3}
    roofline.update(kernel_fractions(k_ms, [g[0] * sd.nsub, g[1], g[2]], (sdom * sd.nsub, sdom, sdom), peak))

    # ---- end to end with HOST buffers: H2D of every local subdomain, step, D2H of every result
    esteps = max(1, min(args.steps, args.e2e_steps))

    def one():
        ctx.h2d(sd.d_a, h_in)
        sd.step(overlap=not args.no_overlap)
        ctx.d2h(h_out, sd.d_a)

    one()
    barrier()
    ctx.timer_start(2)
    for _ in range(esteps):
        one()
    ctx.timer_stop(2)
    ems = ctx.timer_ms(2)
    barrier()
    ems = max_over_ranks(ems) / esteps
    ok = bool(np.isfinite(h_out[512:1024]).all())
    for p in ptrs:
        capi.check(capi.lib().bricks_host_free(p))
    e2e = {"value": pts_step / (ems * 1e-3) * 1e-9, "unit": "GStencil/s", "h2d_bytes_per_step": nbytes,
           "d2h_bytes_per_step": nbytes, "ms_per_step": ems, "steps": esteps, "result_finite": ok,
           "what": "per step and GPU: H2D of all local subdomain storages from pinned host memory, exchange + sweeps, "
                   "D2H of all result storages"}

    if rank == 0:
        line = {"metric": "fp64 GStencil/s", "value": value, "unit": "GStencil/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": make_config(which, world, (dom, sdom)), "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roofline, "overlap": not args.no_overlap,
                "subdomains_rank0": sd.nsub, "exchange_bytes_sent_rank0": sd.exchange_bytes,
                "local_links_rank0": sd.nlinks}
        if world == 1 and not args.no_cpu:
            try:
                line["cpu_baseline"] = cpu_reference_rate(which, budget_s=args.cpu_seconds, dims=(sdom,) * 3)
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": "GStencil/s", "cores": 0, "kind": "reference",
                                        "sample": "unavailable: %r" % (e,)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sd.close()
    return 0


def end_to_end(ctx, sub, field, args, barrier, max_over_ranks, pts_step):
    """Same step, HOST buffers: every step copies its input field host->device from pinned
    memory, runs exchange + sweeps, and copies the resulting field device->host.  The steps are
    software-pipelined (WeakSubdomain.host_steps): upload of step n and download of step n-1
    overlap the sweeps, as a host application streaming fields through the GPU would run them."""
    import ctypes as C
    from bricklib_b200 import capi
    n = sub.dec.nbricks * 512
    ptrs, node = [], C.c_int(-1)
    for _ in range(3):
        p = C.c_void_p()
        if args.e2e_plain_alloc:  # pinned, wherever the pages land (for comparison)
            capi.check(capi.lib().bricks_host_alloc(C.byref(p), n * 8))
        else:  # pinned on the NUMA node of this rank's GPU
            capi.check(capi.lib().bricks_host_alloc_local(ctx._h, C.byref(p), n * 8, C.byref(node)))
        ptrs.append(p)
    h_in, h_out0, h_out1 = (np.ctypeslib.as_array((C.c_double * n).from_address(p.value)) for p in ptrs)
    h_in[:] = field
    steps = max(2, min(args.steps, args.e2e_steps))
    overlap = (not args.no_overlap) and (sub.nranks > 1 or args.overlap)
    # the host link alone, all ranks at once: H2D, then D2H, then both directions together
    nb_own = int(sub.dec.sep_pos[1]) * 512 * 8
    alone = {}
    for name, fn in (("h2d", lambda: ctx.h2d(sub.d_b, h_in, stream=capi.STREAM_H2D, nbytes=nb_own)),
                     ("d2h", lambda: ctx.d2h(h_out0, sub.d_b, stream=capi.STREAM_D2H, nbytes=nb_own)),
                     ("both", lambda: (ctx.h2d(sub.d_b, h_in, stream=capi.STREAM_H2D, nbytes=nb_own),
                                       ctx.d2h(h_out1, sub.d_a, stream=capi.STREAM_D2H, nbytes=nb_own)))):
        fn()
        barrier()
        U, D = capi.STREAM_H2D, capi.STREAM_D2H
        ctx.stream_wait(U, D)
        ctx.timer_start(4, U)          # CUDA events on the copy stream; D2H work is folded in below
        ctx.stream_wait(D, U)
        for _ in range(2):
            fn()
        ctx.stream_wait(U, D)
        ctx.timer_stop(4, U)
        dt = ctx.timer_ms(4) * 1e-3 / 2
        barrier()
        alone[name] = nb_own / max_over_ranks(dt) * 1e-9
    sub.host_steps(h_in, [h_out0, h_out1], 2, overlap)
    barrier()
    ctx.timer_start(2)
    sub.host_steps(h_in, [h_out0, h_out1], steps, overlap)
    ctx.timer_stop(2)
    ms = ctx.timer_ms(2)
    barrier()
    ms = max_over_ranks(ms)
    ok = bool(np.isfinite(h_out0[512:1024]).all() and np.isfinite(h_out1[512:1024]).all()
              and np.array_equal(h_out0[:4096], h_out1[:4096]))
    for p in ptrs:
        capi.check(capi.lib().bricks_host_free(p))
    per_dir = sub.host_step_bytes / (ms / steps * 1e-3) * 1e-9
    link = {"h2d_alone_gbs": alone["h2d"], "d2h_alone_gbs": alone["d2h"], "both_directions_gbs_each": alone["both"],
            "pcie_frac": per_dir / alone["both"], "numa_node_of_pinned_buffers": node.value,
            "pinned_alloc": "bricks_host_alloc" if args.e2e_plain_alloc else "bricks_host_alloc_local",
            "what": "GB/s per direction per GPU with every rank copying at once (max time over ranks); pcie_frac = the "
                    "e2e step's rate per direction / the rate of both directions copying with nothing else running"}
    return {"value": pts_step / (ms / steps * 1e-3) * 1e-9, "unit": "GStencil/s",
            "h2d_bytes_per_step": sub.host_step_bytes, "d2h_bytes_per_step": sub.host_step_bytes,
            "ms_per_step": ms / steps, "steps": steps, "result_finite": ok,
            "host_link_gbs_per_direction_per_gpu": per_dir,
            "host_link": link,
            "what": "per step and GPU: H2D of the input field (all owned bricks; ghost bricks are filled by the "
                    "exchange) from pinned host memory, exchange + sweeps, D2H of the result field (all owned bricks); "
                    "steps software-pipelined over two device buffer sets and two copy streams (upload n+1 and "
                    "download n-1 overlap the sweeps of n)"}


def d4pt9_cpu_reference(dom=64, budget_s=4.0):
    """The reference's generated mpi9pt.py kernel (AVX2, fold 2,2 -- what weak/main-4d.cpp builds on this host;
    oracle/_ref/libref4d.so) under OpenMP on all host threads: full-grid sweeps over a dom^4 subdomain
    (+ ghost bricks), in the build's own fold order as the reference times it."""
    from oracle import oracle4d as D
    if not D.ref_available():
        raise RuntimeError("oracle/_ref/libref4d.so is not built for this host")
    from oracle import refload as R
    threads = use_all_host_threads()
    dec = D.DecompND((dom,) * 4, page_align=False)
    n = dec.nbricks * 256
    a, b = R.aligned_empty(n), R.aligned_empty(n)
    a[:] = synth_uniform(n, 1)
    b[:] = 0.0
    blist = D.sweep_list(dec.grid, 0)
    t1 = D.ref_brick_sweep_timed(dec.adj, blist, a, b, 256, 2)
    reps = int(max(2, min(200, budget_s / max(t1, 1e-4))))
    reps += reps & 1
    t = D.ref_brick_sweep_timed(dec.adj, blist, a, b, 256, reps)
    pts = float(len(blist)) * 256
    return {"full_grid_sweep_ms": t * 1e3, "gstencil_per_s_full_grid": pts / t * 1e-9, "cores": threads,
            "kind": "reference",
            "sample": "%d full-grid sweeps over one %d^4 subdomain (+ghost bricks: %d bricks), reference codegen "
                      "kernel of stencils/mpi9pt.py (AVX2, fold 2,2) under OpenMP with %d threads" %
                      (reps, dom, len(blist), threads)}


def d4pt9_block(peak, dom=64, iters=10):
    """The 4-D path (weak/main-4d.cpp): drivers/weak4d -s 64 -v in a process of its own
    (plain C++ above the C ABI, no Python): one 64^4 subdomain (+ ghost: 20^4 bricks of 4x4x4x4, 328 MB per
    field), validated against the host array twin, then exchange + 8 shrinking sweeps timed with CUDA events.
    An extra block: a failure here is reported, it does not take the headline line down."""
    import re
    import subprocess
    exe = os.path.join(ROOT, "drivers", "_build", "weak4d")
    try:
        r = subprocess.run([exe, "-s", str(dom), "-I", str(iters), "-v"], capture_output=True, text=True, timeout=180)
        out = r.stdout
        m_full = re.search(r"full-grid sweep \(ms\): ([0-9.e+-]+); HBM \(GB/s per GPU, 16 B per point\): ([0-9.e+-]+)", out)
        m_perf = re.search(r"perf ([0-9.e+-]+) GStencil/s", out)
        m_par = re.search(r"(result match|result mismatch): max relative error vs the host array twin ([0-9.e+-]+)", out)
        if r.returncode != 0 or not (m_full and m_perf and m_par):
            return {"error": "drivers/weak4d rc %d" % r.returncode, "tail": (out + r.stderr)[-400:]}
        gbs = float(m_full.group(2))
        try:
            cpu = d4pt9_cpu_reference(dom)
            cpu["gpu_over_cpu_full_grid_sweep"] = cpu["full_grid_sweep_ms"] / float(m_full.group(1))
        except Exception as e:  # noqa: BLE001
            cpu = {"error": repr(e)}
        return {"cpu_baseline": cpu, "workload": "d4pt9: %d^4 subdomain per GPU, 4x4x4x4 bricks, ghost depth 8, 9-point star, exchange + 8 "
                            "sweeps over shrinking regions (weak/main-4d.cpp)" % dom,
                "parity": {"ok": m_par.group(1) == "result match", "max_relerr_vs_host_array_twin": float(m_par.group(2)),
                           "bar": 1e-12},
                "gstencil_per_s": float(m_perf.group(1)), "full_grid_sweep_ms": float(m_full.group(1)),
                "achieved_gbs": gbs, "frac_of_measured_peak": gbs / peak if peak else None,
                # SURVEY.md 8d counts interior points only: with ghost depth 8 around 64^4 the period's eight
                # shrinking sweeps update 1.95 x the interior points (4 x 20^4 + 3 x 18^4 + 16^4 bricks)
                "period_frac_interior": float(m_perf.group(1)) * 16.0 / peak if peak else None,
                "definition": "frac_of_measured_peak = 16 B x every point of one full-grid sweep (20^4 bricks) / its "
                              "time / peak; period_frac_interior = 16 B x interior GStencil/s of the whole period "
                              "(exchange + 8 sweeps) / peak",
                "algorithmic_bytes_per_point": 16, "kernel": "line_kernel (bricklib_b200/csrc/stencil4d.cu)"}
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)}


def all_stencils(ctx, sub, peak):
    """Kernel-only numbers for every stencil of the path on the same 66^3-brick grid."""
    out = {}
    for w in (7, 13, 25, 125):
        sub.upload(sub.host_field())
        ms = kernel_probe(ctx, sub, w, reps=5)
        fr = kernel_fractions(ms, sub.gdims, SUBDOMAIN, peak)
        out[str(w)] = {"ms_per_sweep": ms, "achieved_gbs": fr["achieved"], "frac_of_measured_peak": fr["frac"],
                       "frac_interior": fr["frac_interior"],
                       "frac_interior_of_8TBps_nominal": fr["frac_interior_of_8TBps_nominal"],
                       "interior_gstencil_per_s": fr["interior_gstencil_per_s_kernel_only"]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="weak125", choices=sorted(WORKLOADS))
    ap.add_argument("--no-overlap", action="store_true", help="serial exchange then sweeps (reference schedule); the "
                    "default on ONE GPU, where the exchange is a 0.04 ms on-device copy and the inner + shell split of the "
                    "first sweep costs more than it hides (2.453 vs 2.354 ms/step, profiles/r03_scaling/diag_n1)")
    ap.add_argument("--overlap", action="store_true", help="force the overlapped schedule on one GPU too")
    ap.add_argument("--no-early", action="store_true", help="start each exchange at the step boundary only")
    ap.add_argument("--shift-exchange", action="store_true",
                    help="weak workloads: the SHIFT exchange (3 stages, 6 messages) instead of one message per region")
    ap.add_argument("--exchange", default=None, choices=["regions", "packed", "shift"],
                    help="weak workloads: one send/recv pair per region (default), one packed message per peer, or SHIFT")
    ap.add_argument("--unsplit", action="store_true", default=False,
                    help="with the early exchange: run a first sweep whose exchange is already in flight as ONE launch "
                         "(+1.5 %% at 2 GPUs, -3.5 %% at 8 where the exchange outlasts the sweep it hides behind: "
                         "profiles/r03_scaling/)")
    ap.add_argument("--no-unsplit", dest="unsplit", action="store_false",
                    help="keep the inner + shell split of every first sweep (default)")
    ap.add_argument("--p2p", action="store_true", default=True,
                    help="N > 1: the copy-engine exchange (cudaMemcpyAsync into peer-mapped ghost bricks; default)")
    ap.add_argument("--no-p2p", dest="p2p", action="store_false", help="N > 1: NCCL send/recv instead")
    ap.add_argument("--early", action="store_true", help="start each exchange behind the previous step's last sweep "
                    "(default with NCCL, whose exchange outlasts the sweep it hides behind; off with the copy engines)")
    ap.add_argument("--trace", default=None, metavar="FILE",
                    help="write every rank's timeline of one traced step (bricks_step_trace) to FILE.rank<r>.json")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--e2e-plain-alloc", action="store_true",
                    help="e2e: pinned host buffers from bricks_host_alloc (no NUMA placement), for comparison")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity block in front of the timed loop")
    ap.add_argument("--no-all-stencils", action="store_true", help="skip the kernel-only numbers for 7/13/25/125-point")
    ap.add_argument("--all-stencils", action="store_true", help="(the default now; kept for old command lines)")
    ap.add_argument("--no-4d", action="store_true", help="skip the 4-D block (drivers/weak4d on a 64^4 subdomain, N = 1 only)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling block (1024^3 in 128^3, 125-point)")
    ap.add_argument("--dom", type=int, default=STRONG_DOM, help="strong workloads: global grid edge")
    ap.add_argument("--sdom", type=int, default=STRONG_SDOM, help="strong workloads: subdomain edge")
    args = ap.parse_args()
    args.launcher_omp = os.environ.get("OMP_NUM_THREADS", "unset")
    if args.exchange is None:
        args.exchange = "shift" if args.shift_exchange else "regions"
    args.shift_exchange = args.exchange == "shift"
    which = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference(args, which)
    if args.workload.startswith("strong"):
        return run_gpu_strong(args, which)
    return run_gpu(args, which)


if __name__ == "__main__":
    sys.exit(main())
