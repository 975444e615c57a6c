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
import numpy as np

from . import capi


def shell_boxes(g, inner_lo, inner_hi):
    """Six disjoint slabs covering [0,g)^3 minus [inner_lo, inner_hi)^3 (i fastest)."""
    (g0, g1, g2), (a0, a1, a2), (b0, b1, b2) = g, inner_lo, inner_hi
    return [((0, 0, 0), (g0, g1, a2)), ((0, 0, b2), (g0, g1, g2)),
            ((0, 0, a2), (g0, a1, b2)), ((0, b1, a2), (g0, g1, b2)),
            ((0, a1, a2), (a0, b1, b2)), ((b0, a1, a2), (g0, b1, b2))]


def shell_between(outer_lo, outer_hi, inner_lo, inner_hi):
    """Six disjoint slabs covering [outer_lo, outer_hi) minus [inner_lo, inner_hi) (i fastest)."""
    (o0, o1, o2), (g0, g1, g2) = outer_lo, outer_hi
    (a0, a1, a2), (b0, b1, b2) = inner_lo, inner_hi
    return [((o0, o1, o2), (g0, g1, a2)), ((o0, o1, b2), (g0, g1, g2)),
            ((o0, o1, a2), (g0, a1, b2)), ((o0, b1, a2), (g0, g1, b2)),
            ((o0, a1, a2), (a0, b1, b2)), ((b0, a1, a2), (g0, b1, b2))]


class WeakSubdomain(object):
    """One rank's subdomain: BrickDecomp tables, two ping-pong storages on the device, the box
    plans of the sweep schedule and the exchange plan."""

    def __init__(self, ctx, dims, stencil, rank=0, procdims=(1, 1, 1), depth=8, seed=20240601, shift=False):
        self.ctx, self.stencil, self.rank = ctx, stencil, rank
        self.procdims = list(procdims)
        self.nranks = int(np.prod(procdims))
        self.dims = tuple(dims)
        self.dec = capi.Decomp(dims, depth)
        self.st_iter = capi.ST_ITER[stencil]
        self.coo = capi.cart_coords(rank, self.procdims)
        self.rank_map = self.dec.rank_map(self.procdims, self.coo)
        g = self.dec.gdims
        gb = depth // 8
        self.gdims = g
        nb = self.dec.nbricks
        self.nbytes = nb * 512 * 8
        self.d_a = ctx.malloc(self.nbytes)
        self.d_b = ctx.malloc(self.nbytes)
        ctx.memset(self.d_a, 0, self.nbytes)
        ctx.memset(self.d_b, 0, self.nbytes)
        full = ((0, 0, 0), tuple(g))
        inner = (tuple(2 * gb for _ in g), tuple(x - 2 * gb for x in g))
        owned = (tuple(gb for _ in g), tuple(x - gb for x in g))
        self.plan_full = capi.Plan(ctx, self.dec.grid, self.dec.adj, [full])
        self.plan_inner = capi.Plan(ctx, self.dec.grid, self.dec.adj, [inner])
        self.plan_shell = capi.Plan(ctx, self.dec.grid, self.dec.adj, shell_boxes(g, inner[0], inner[1]))
        self.plan_owned = capi.Plan(ctx, self.dec.grid, self.dec.adj, [owned])
        # the owned bricks that some neighbour needs (skin) = owned box minus the inner box
        self.plan_oskin = capi.Plan(ctx, self.dec.grid, self.dec.adj,
                                    shell_between(owned[0], owned[1], inner[0], inner[1]))
        self._exchange_in_flight = None  # storage whose exchange was started by the previous step
        self._pairs = None               # device storage pairs of the host-fed pipeline (host_steps)
        self.host_step_bytes = int(self.dec.sep_pos[1]) * 512 * 8
        ro, so, nbts, rp, sp = self.dec.exchange_plan(512)
        self.exchange_bytes = int(nbts.sum())
        if shift:
            # SHIFT exchange (multiStageExchangeView, brick-mpi.h:587-681): 3 stages, 6 messages
            self.exchange = capi.ShiftExchange(ctx, rank, self.dec.shift_plan(512))
        else:
            self.exchange = capi.Exchange(ctx, rank, ro, so, nbts, rp, sp)
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
        self._exchange_in_flight = None
        self.ctx.h2d(self.d_a, field)
        self.ctx.sync()

    def download(self):
        return self.ctx.download(self.d_a, self.dec.nbricks * 512)

    # ------------------------------------------------------------------ one step
    def step(self, overlap=True, coeff=None, pair=None, early=False):
        """exchange(d_a) + ST_ITER ping-pong sweeps; the result ends in d_a (ST_ITER is even).
        Everything is enqueued asynchronously; returns the number of kernels launched.
        `pair` = (in/out storage, scratch storage) overrides (d_a, d_b).
        `early`: the last sweep computes the skin bricks first and the exchange of the NEXT step
        starts as soon as they are written, so it also hides behind the rest of that sweep (only
        valid when the next step continues from this result; any host access to the field must
        be preceded by ctx.sync(), which covers the communication stream too)."""
        C, X = capi.STREAM_COMPUTE, capi.STREAM_COMM
        n0 = capi.launch_count()
        ctx, st = self.ctx, self.stencil
        d_a, d_b = pair if pair is not None else (self.d_a, self.d_b)
        if self._exchange_in_flight != d_a:
            # the exchange reads the skin written by the previous step's last sweep
            ctx.stream_wait(X, C)
            self.exchange.run(d_a, X)
        self._exchange_in_flight = None
        src, dst = d_a, d_b
        for it in range(self.st_iter):
            last = it == self.st_iter - 1
            if it == 0:
                if overlap:
                    self.plan_inner.stencil(st, src, dst, 512, 512, coeff, C)
                    ctx.stream_wait(C, X)
                    self.plan_shell.stencil(st, src, dst, 512, 512, coeff, C)
                else:
                    ctx.stream_wait(C, X)
                    self.plan_full.stencil(st, src, dst, 512, 512, coeff, C)
            elif last and early and overlap:
                self.plan_oskin.stencil(st, src, dst, 512, 512, coeff, C)
                ctx.stream_wait(X, C)
                self.exchange.run(dst, X)  # dst == d_a: input of the next step
                self._exchange_in_flight = dst
                self.plan_inner.stencil(st, src, dst, 512, 512, coeff, C)
            elif last:
                self.plan_owned.stencil(st, src, dst, 512, 512, coeff, C)
            else:
                self.plan_full.stencil(st, src, dst, 512, 512, coeff, C)
            src, dst = dst, src
        assert src == d_a
        # NCCL launches its own kernels; count ours plus one per remote exchange
        self.launches_per_step = capi.launch_count() - n0
        return self.launches_per_step

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
        if self._exchange_in_flight is not None:  # an early exchange of a device-resident run
            ctx.sync()
            self._exchange_in_flight = None
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
        for p in (self.plan_full, self.plan_inner, self.plan_shell, self.plan_owned, self.plan_oskin):
            p.close()
        self.exchange.close()
        for a1, b1 in (self._pairs or [])[1:]:
            self.ctx.free(a1)
            self.ctx.free(b1)
        self.ctx.free(self.d_a)
        self.ctx.free(self.d_b)
        self.dec.close()


def interior_points(dims, nranks=1):
    """tot_elems of the reference (weak/args.cpp:107-130): global interior points per sweep."""
    return int(np.prod(dims)) * nranks
