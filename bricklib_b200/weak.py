"""Weak-scaling step: one subdomain per GPU, one halo exchange per ST_ITER stencil sweeps.

Host-side mirror of the reference's weak driver hot loop (weak/main.cu:246-287 `brick_func`;
CPU twin weak/main.cpp:172-213) above the C ABI:

    reference                                   here
    ---------                                   ----
    cudaMemcpy D2H skin / MPI / H2D ghost       ExchangeView.start(): NCCL send/recv (or on-device
    or CUDA-aware bDecomp.exchange()            copies for self neighbours) on the comm stream
    brick_kernel<<<strideb^3,32>>> x ST_ITER    bricks_stencil over box plans on the compute stream

and two things the reference does not do: (1) the first sweep of a step runs on the bricks that
do not read ghost data ([2, n-2) cells of the brick grid = the inner-inner region
[1, sep_pos[0]) of the numbering) WHILE the exchange is in flight, and only its outer shell
waits for the exchange; (2) the last sweep skips the ghost bricks (the CPU driver's skip=1,
weak/main.cpp:204-209): their output is never read because the next exchange overwrites it.
"""
import os as _os

import numpy as np

from . import capi


class WeakSubdomain(object):
    """One rank's subdomain: BrickDecomp tables, two ping-pong storages on the device and the
    library's step object (bricks_step_*: the box plans of the sweep schedule over one device copy
    of grid / adj, and the communication program).

    exchange: "regions" -- one ncclSend/ncclRecv pair per BrickDecomp region straight out of the
                           storage (BrickDecomp::exchange, brick-mpi.h:466-495; 42 pairs at depth 8)
              "packed"  -- one message per peer rank (the role of ExchangeView, brick-mpi.h:96-123,
                           525-580: one message per neighbour; the mmap views become a pack kernel,
                           one send/recv per peer and an unpack kernel over a device arena)
              "shift"   -- MultiStageExchangeView (brick-mpi.h:587-681): 3 stages, 6 messages"""

    def __init__(self, ctx, dims, stencil, rank=0, procdims=(1, 1, 1), depth=8, seed=20240601, shift=False,
                 exchange=None):
        self.ctx, self.stencil, self.rank = ctx, stencil, rank
        self.procdims = list(procdims)
        self.nranks = int(np.prod(procdims))
        self.dims = tuple(dims)
        self.dec = capi.Decomp(dims, depth)
        self.st_iter = capi.ST_ITER[stencil]
        self.coo = capi.cart_coords(rank, self.procdims)
        self.rank_map = self.dec.rank_map(self.procdims, self.coo)
        g = self.dec.gdims
        self.gdims = g
        nb = self.dec.nbricks
        self.nbytes = nb * 512 * 8
        self.d_a = ctx.malloc(self.nbytes)
        self.d_b = ctx.malloc(self.nbytes)
        ctx.memset(self.d_a, 0, self.nbytes)
        ctx.memset(self.d_b, 0, self.nbytes)
        ro, so, nbts, rp, sp = self.dec.exchange_plan(512)
        self.exchange_bytes = int(nbts.sum())
        kind = exchange or ("shift" if shift else "regions")
        # the schedule and its communication program are built by the C++ host layer
        # (BrickDecomp::makeStep) and run by the library (csrc/step.cu)
        self.step_obj, self.exchange_kind = self.dec.make_step(ctx, rank, self.nranks, kind, 512)
        self.plan_full = self.step_obj.plan(capi.PLAN_FULL)
        self.plan_inner = self.step_obj.plan(capi.PLAN_INNER)
        self.plan_shell = self.step_obj.plan(capi.PLAN_SHELL)
        self.plan_owned = self.step_obj.plan(capi.PLAN_OWNED)
        self.plan_oskin = self.step_obj.plan(capi.PLAN_OSKIN)
        self._pairs = None               # device storage pairs of the host-fed pipeline (host_steps)
        self.host_step_bytes = int(self.dec.sep_pos[1]) * 512 * 8
        self.seed = seed + 1000 * rank
        self.launches_per_step = None

    # ------------------------------------------------------------------ data
    def host_field(self):
        """U[0,1) initial field over every brick of the storage (ghost bricks included; the
        first exchange overwrites them).  Brick 0, the sentinel, is zero."""
        f = capi.uniform(self.dec.nbricks * 512, seed=self.seed)
        f[:512] = 0.0
        return f

    def upload(self, field):
        self.ctx.sync()
        self.step_obj.reset()
        self.ctx.h2d(self.d_a, field)
        self.ctx.sync()

    def download(self):
        return self.ctx.download(self.d_a, self.dec.nbricks * 512)

    # ------------------------------------------------------------------ one step
    def step(self, overlap=True, coeff=None, pair=None, early=False, unsplit=None):
        """exchange(d_a) + ST_ITER ping-pong sweeps; the result ends in d_a (ST_ITER is even).
        The schedule itself lives in the library (bricks_step_run, csrc/step.cu): everything is
        enqueued asynchronously; returns the number of kernels launched.
        `pair` = (in/out storage, scratch storage) overrides (d_a, d_b).
        `early`: the last sweep computes the skin bricks first and the exchange of the NEXT step
        starts as soon as they are written, so it also hides behind the rest of that sweep (only
        valid when the next step continues from this result; any host access to the field must
        be preceded by ctx.sync(), which covers the communication stream too).
        `unsplit`: with early, a first sweep whose exchange is already in flight runs as one launch
        (default: BRICKS_EARLY_UNSPLIT=1 in the environment)."""
        n0 = capi.launch_count()
        d_a, d_b = pair if pair is not None else (self.d_a, self.d_b)
        flags = (capi.STEP_OVERLAP if overlap else 0) | (capi.STEP_EARLY if early else 0)
        if unsplit is None:
            unsplit = _os.environ.get("BRICKS_EARLY_UNSPLIT", "0") == "1"
        if unsplit:
            flags |= capi.STEP_EARLY_UNSPLIT
        self.step_obj.run(self.stencil, self.st_iter, d_a, d_b, coeff, flags)
        # NCCL launches its own kernels; count ours plus one per remote exchange
        self.launches_per_step = capi.launch_count() - n0
        return self.launches_per_step

    def enable_p2p(self, allgather):
        """Peer-memory exchange (bricks_step_p2p_*): the region messages of d_a are stored straight
        into the peers' ghost bricks over NVLink instead of going through NCCL send/recv.
        `allgather(bytes)` returns every rank's bytes in rank order (collective: call on all ranks).
        All-or-nothing: if export or connect fails on ANY rank (CUDA IPC disabled in a container, no
        peer access, no stream memory operations) every rank goes back to NCCL and this returns False
        with the reason in self.p2p_error.  Storages other than d_a (the second pair of host_steps)
        keep the NCCL path."""
        self.p2p_error = None
        desc = b"\0" * 256
        try:
            desc = self.step_obj.p2p_export(self.d_a, self.rank)
        except capi.BricksError as e:
            self.p2p_error = str(e)
        descs = allgather(desc + (b"\1" if self.p2p_error is None else b"\0"))
        if all(d[256:] == b"\1" for d in descs):
            try:
                self.step_obj.p2p_connect([d[:256] for d in descs])
            except capi.BricksError as e:
                self.p2p_error = str(e)
        elif self.p2p_error is None:
            self.p2p_error = "a peer could not export its storage"
        oks = allgather(b"\1" if self.p2p_error is None else b"\0")
        if not all(o == b"\1" for o in oks):
            if self.p2p_error is None:
                self.p2p_error = "a peer could not map this rank's storage"
            self.step_obj.p2p_disable()
            return False
        self.exchange_kind = (self.exchange_kind.replace("one ncclSend/ncclRecv pair per region", "one message per region")
                              + ", stored straight into peer-mapped ghost bricks over NVLink (no NCCL kernel)")
        return True

    def exchange_only(self, storage=None):
        """the communication program alone, on the communication stream"""
        self.step_obj.exchange(self.d_a if storage is None else storage)

    def sweep_full(self, stencil, src, dst, coeff=None):
        """one sweep over every brick of the grid (ghost bricks included): the dominant kernel"""
        self.step_obj.sweep(capi.PLAN_FULL, stencil, src, dst, coeff)

    # ------------------------------------------------------------------ host-fed steps
    def host_steps(self, h_in, h_outs, steps, overlap=True):
        """`steps` independent steps fed from and drained to HOST memory (pinned numpy views):
        step n uploads h_in, runs exchange + sweeps and downloads the result to h_outs[n % 2].
        Software-pipelined over two device storage pairs and the two copy streams: the upload
        of step n and the download of step n-1 run while step n-1 / n computes, so a step costs
        max(upload, download, compute) instead of their sum.  Returns after everything is
        enqueued; the caller synchronises."""
        ctx = self.ctx
        C, U, D = capi.STREAM_COMPUTE, capi.STREAM_H2D, capi.STREAM_D2H
        ctx.sync()  # an early exchange of a device-resident run may still be in flight
        self.step_obj.reset()
        # ghost bricks are neither input (the exchange fills them) nor result: BrickDecomp numbers
        # them last, so only bricks [0, sep_pos[1]) cross the bus
        nbytes = int(self.dec.sep_pos[1]) * 512 * 8
        self.host_step_bytes = nbytes
        if self._pairs is None:
            a1, b1 = ctx.malloc(self.nbytes), ctx.malloc(self.nbytes)
            ctx.memset(b1, 0, self.nbytes)
            ctx.memset(a1, 0, self.nbytes)
            self._pairs = [(self.d_a, self.d_b), (a1, b1)]
        for n in range(steps + 1):
            if n < steps:
                pa = self._pairs[n & 1]
                ctx.stream_wait(U, D)          # the download of step n-2 has drained this pair
                ctx.h2d(pa[0], h_in, stream=U, nbytes=nbytes)
            if n >= 1:
                pb = self._pairs[(n - 1) & 1]
                ctx.stream_wait(D, C)          # step n-1 has finished computing
                ctx.d2h(h_outs[(n - 1) & 1], pb[0], stream=D, nbytes=nbytes)
            if n < steps:
                ctx.stream_wait(C, U)          # the upload of step n has landed
                self.step(overlap=overlap, pair=pa)
        ctx.stream_wait(C, D)

    def close(self):
        self.step_obj.close()
        for a1, b1 in (self._pairs or [])[1:]:
            self.ctx.free(a1)
            self.ctx.free(b1)
        self.ctx.free(self.d_a)
        self.ctx.free(self.d_b)
        self.dec.close()


def interior_points(dims, nranks=1):
    """tot_elems of the reference (weak/args.cpp:107-130): global interior points per sweep."""
    return int(np.prod(dims)) * nranks
