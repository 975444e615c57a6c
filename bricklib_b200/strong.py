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

        # the schedule and its communication program live in the library (csrc/step.cu)
        S = self.step_obj = capi.Step(ctx, self.dec.grid, self.dec.adj, depth // 8)
        self.plan_full = S.plan(capi.PLAN_FULL)
        self.plan_owned = S.plan(capi.PLAN_OWNED)
        self.plan_inner = S.plan(capi.PLAN_INNER)
        self.plan_shell = S.plan(capi.PLAN_SHELL)
        self.has_inner = self.plan_inner is not None

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
        pack, unpack = [], []
        for k, r in enumerate(peers):
            if r in send:
                pack += [(self.in_a[s] + o, so[k] + bo, n) for s, _, o, n, bo, _ in send[r]["pieces"]]
            if r in recv:
                unpack += [(ro[k] + bo, self.in_a[s] + o, n) for s, _, o, n, bo, _ in recv[r]["pieces"]]
        # communication program: pack -> per-peer messages -> unpack -> intra-GPU links
        # (strong/main.cu:416-435 and 201-208: neighbour's skin[i] -> my ghost[i])
        if peers:
            S.arena(max(off, 256))
            S.add_copy(pack, capi.BASE_ABS, capi.BASE_ARENA)
            S.add_exchange(capi.ExchangeV(ctx, rank, ro, rb, peers, so, sb, peers), capi.BASE_ARENA)
            S.add_copy(unpack, capi.BASE_ARENA, capi.BASE_ABS)
        self.nlinks = len(self.plan.links)
        S.add_copy([(self.in_a[ss] + so_, self.in_a[ds] + do, n) for ss, ds, so_, do, n in self.plan.links],
                   capi.BASE_ABS, capi.BASE_ABS)

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
    def exchange_only(self):
        """pack -> per-peer messages -> unpack -> local links, on the communication stream."""
        self.step_obj.exchange(self.in_a[0])

    def sweep_full(self, stencil, src, dst, coeff=None):
        """one launch over every brick of every local subdomain"""
        self.step_obj.sweep(capi.PLAN_FULL, stencil, src, dst, coeff)

    def step(self, overlap=True, coeff=None):
        """bricks_step_run over all local subdomains: each sweep is ONE launch batched over them"""
        n0 = capi.launch_count()
        self.step_obj.run(self.stencil, self.st_iter, self.in_a, self.in_b, coeff,
                          capi.STEP_OVERLAP if overlap else 0)
        return capi.launch_count() - n0

    def close(self):
        self.step_obj.close()
        for d in (self.d_a, self.d_b):
            self.ctx.free(d)
        self.plan.close()
        self.dec.close()
