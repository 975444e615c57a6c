"""Strong-scaling step: a fixed dom^3 domain cut into (dom/sdom)^3 subdomains, several per GPU.

Host-side mirror of the reference's strong driver hot loop (strong/main.cu:397-510 `brick_func`)
above the C ABI.  Ownership, Morton neighbours and the link / message lists come from the C++
planner (include/bricks/strong-plan.h); this module only allocates, wires pointers and enqueues:

    reference                                         here
    ---------                                         ----
    cudaCopy<<<nlinks,64>>>(local links)              bricks_copy_regions (LinkTable local)
    MPI_Irecv/Isend one packed buffer per peer        bricks_exchange_create_v over one arena (NCCL)
    cudaCopy unpack / pack                            LinkTable unpack / pack
    brick_kernel<<<(STRIDEB*nsub,STRIDEB,STRIDEB),32>>> x ST_ITER
                                                      ONE bricks_stencil launch per sweep, batched
                                                      over all local subdomains (blockIdx.y)

As in the weak step, the first sweep's inner box (cells [2, n-2) of every subdomain, which reads
no ghost brick) runs while pack -> exchange -> unpack -> local links are in flight on the
communication stream, and the last sweep skips the ghost bricks.
"""
import numpy as np

from . import capi
from .weak import shell_boxes


class StrongDomain(object):
    def __init__(self, ctx, dom, sdom, stencil, rank=0, nranks=1, depth=8, seed=20240601):
        self.ctx, self.stencil, self.rank, self.nranks = ctx, stencil, rank, nranks
        self.dom, self.sdom = dom, sdom
        self.dec = capi.Decomp((sdom, sdom, sdom), depth)
        self.plan = capi.StrongPlan(dom, sdom, nranks, rank, self.dec)
        self.nsub = self.plan.nsub
        self.st_iter = capi.ST_ITER[stencil]
        self.seed = seed
        nb = self.dec.nbricks
        self.sub_bytes = nb * 512 * 8
        # one arena per field: subdomain s lives at base + s * sub_bytes
        self.d_a = ctx.malloc(self.nsub * self.sub_bytes)
        self.d_b = ctx.malloc(self.nsub * self.sub_bytes)
        ctx.memset(self.d_a, 0, self.nsub * self.sub_bytes)
        ctx.memset(self.d_b, 0, self.nsub * self.sub_bytes)
        self.in_a = [self.d_a + s * self.sub_bytes for s in range(self.nsub)]
        self.in_b = [self.d_b + s * self.sub_bytes for s in range(self.nsub)]

        g = self.dec.gdims
        gb = depth // 8
        full = ((0, 0, 0), tuple(g))
        inner = (tuple(2 * gb for _ in g), tuple(x - 2 * gb for x in g))
        owned = (tuple(gb for _ in g), tuple(x - gb for x in g))
        P = capi.Plan
        self.plan_full = P(ctx, self.dec.grid, self.dec.adj, [full])
        self.plan_owned = P(ctx, self.dec.grid, self.dec.adj, [owned])
        self.has_inner = all(b > a for a, b in zip(inner[0], inner[1]))
        if self.has_inner:
            self.plan_inner = P(ctx, self.dec.grid, self.dec.adj, [inner])
            self.plan_shell = P(ctx, self.dec.grid, self.dec.adj, shell_boxes(g, inner[0], inner[1]))

        # intra-GPU neighbours: neighbour's skin[i] -> my ghost[i] (strong/main.cu:201-208)
        self.local = capi.LinkTable(ctx, [(self.in_a[ss] + so, self.in_a[ds] + do, n)
                                          for ss, ds, so, do, n in self.plan.links])
        # inter-GPU: one packed buffer per peer and direction, all inside one arena
        peers = sorted(set(p["rank"] for p in self.plan.send) | set(p["rank"] for p in self.plan.recv))
        send = {p["rank"]: p for p in self.plan.send}
        recv = {p["rank"]: p for p in self.plan.recv}
        off, so, sb, ro, rb = 0, [], [], [], []
        for r in peers:
            n = send[r]["bytes"] if r in send else 0
            so.append(off), sb.append(n)
            off += (n + 255) // 256 * 256
        for r in peers:
            n = recv[r]["bytes"] if r in recv else 0
            ro.append(off), rb.append(n)
            off += (n + 255) // 256 * 256
        self.arena_bytes = off
        self.exchange_bytes = int(sum(sb))
        self.arena = ctx.malloc(max(off, 256))
        pack, unpack = [], []
        for k, r in enumerate(peers):
            if r in send:
                pack += [(self.in_a[s] + o, self.arena + so[k] + bo, n) for s, _, o, n, bo, _ in send[r]["pieces"]]
            if r in recv:
                unpack += [(self.arena + ro[k] + bo, self.in_a[s] + o, n) for s, _, o, n, bo, _ in recv[r]["pieces"]]
        self.pack, self.unpack = capi.LinkTable(ctx, pack), capi.LinkTable(ctx, unpack)
        self.exchange = capi.ExchangeV(ctx, rank, ro, rb, peers, so, sb, peers) if peers else None

    # ------------------------------------------------------------------ data
    def host_field(self, local_index):
        """U[0,1) field of local subdomain `local_index` (seeded by its Morton id); brick 0 zero."""
        f = capi.uniform(self.dec.nbricks * 512, seed=self.seed + self.plan.first_id + local_index)
        f[:512] = 0.0
        return f

    def upload(self, fields=None):
        for s in range(self.nsub):
            f = self.host_field(s) if fields is None else fields[s]
            self.ctx.h2d(self.in_a[s], f)
            self.ctx.sync()

    def download(self, local_index):
        return self.ctx.download(self.in_a[local_index], self.dec.nbricks * 512)

    # ------------------------------------------------------------------ one step
    def exchange_only(self, stream=capi.STREAM_COMM):
        """pack -> per-peer messages -> unpack -> local links, all on one stream."""
        self.pack.run(stream)
        if self.exchange is not None:
            self.exchange.run(self.arena, stream)
        self.unpack.run(stream)
        self.local.run(stream)

    def step(self, overlap=True, coeff=None):
        C_, X = capi.STREAM_COMPUTE, capi.STREAM_COMM
        ctx, st = self.ctx, self.stencil
        n0 = capi.launch_count()
        ctx.stream_wait(X, C_)
        self.exchange_only(X)
        src, dst = self.in_a, self.in_b
        for it in range(self.st_iter):
            last = it == self.st_iter - 1
            if it == 0:
                if overlap and self.has_inner:
                    self.plan_inner.stencil(st, src, dst, 512, 512, coeff, C_)
                    ctx.stream_wait(C_, X)
                    self.plan_shell.stencil(st, src, dst, 512, 512, coeff, C_)
                else:
                    ctx.stream_wait(C_, X)
                    self.plan_full.stencil(st, src, dst, 512, 512, coeff, C_)
            elif last:
                self.plan_owned.stencil(st, src, dst, 512, 512, coeff, C_)
            else:
                self.plan_full.stencil(st, src, dst, 512, 512, coeff, C_)
            src, dst = dst, src
        assert src is self.in_a
        return capi.launch_count() - n0

    def close(self):
        plans = [self.plan_full, self.plan_owned] + ([self.plan_inner, self.plan_shell] if self.has_inner else [])
        for p in plans:
            p.close()
        if self.exchange is not None:
            self.exchange.close()
        for t in (self.local, self.pack, self.unpack):
            t.close()
        for d in (self.d_a, self.d_b, self.arena):
            self.ctx.free(d)
        self.plan.close()
        self.dec.close()
