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

        # the schedule and its communication program (pack -> one message per peer -> unpack ->
        # intra-GPU links; strong/main.cu:397-447, 201-208) are built by the C++ host layer
        # (makeStrongStep, include/bricks/strong-plan.h) and run by the library (csrc/step.cu)
        h, sent = capi.C.c_void_p(), capi.C.c_size_t()
        capi.check(capi.lib().bricks_strong_step_create(self.plan._h, ctx._h, rank, self.dec._h,
                                                        capi.C.c_void_p(self.d_a), self.sub_bytes,
                                                        capi.C.byref(h), capi.C.byref(sent)))
        S = self.step_obj = capi.Step(ctx, handle=h.value)
        self.exchange_bytes = int(sent.value)
        self.nlinks = len(self.plan.links)
        self.plan_full = S.plan(capi.PLAN_FULL)
        self.plan_owned = S.plan(capi.PLAN_OWNED)
        self.plan_inner = S.plan(capi.PLAN_INNER)
        self.plan_shell = S.plan(capi.PLAN_SHELL)
        self.has_inner = self.plan_inner is not None

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
