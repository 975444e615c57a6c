"""Halo exchange, region copies and the weak-scaling step on the GPU, bit-exact (byte work) or
to 1e-12 (after stencil sweeps) against the oracle's restatement of BrickDecomp::exchange
(include/brick-mpi.h:466-495) and against an independent periodic array computation."""
import threading

import numpy as np
import pytest

from conftest import REL_TOL, relerr

pytestmark = pytest.mark.gpu


def test_copy_regions_is_bit_exact(ctx, built):
    """bricks_copy_regions (cudaCopy of strong/main.cu:76-83): ragged sizes, many links."""
    rng = np.random.default_rng(0)
    n = 1 << 20
    src = rng.integers(0, 2 ** 62, n, dtype=np.int64)
    d_src = ctx.upload(src)
    d_dst = ctx.upload(np.zeros(n, np.int64))
    links, exp = [], np.zeros(n, np.int64)
    pos = 0
    for nb16 in [1, 2, 3, 255, 256, 257, 4096, 4097, 70000, 1, 9000]:
        ne = nb16 * 2  # int64 elements = 16-byte units * 2
        s0 = int(rng.integers(0, n - ne) // 2 * 2)
        links.append((d_src + s0 * 8, d_dst + pos * 8, ne * 8))
        exp[pos:pos + ne] = src[s0:s0 + ne]
        pos += ne + 2 * int(rng.integers(0, 5))
    built.copy_regions(ctx, links)
    ctx.sync()
    got = ctx.download(d_dst, n, np.int64)
    assert np.array_equal(got, exp)
    ctx.free(d_src)
    ctx.free(d_dst)


@pytest.mark.parametrize("dims", [(16, 16, 16), (32, 24, 16), (64, 64, 64)])
def test_self_exchange_matches_oracle_bit_exact(ctx, built, oracle, dims):
    """np=1 on a periodic grid: every neighbour is this rank (docs/testcases.md:45-46 of the
    reference runs exactly this), so all 42 regions become on-device skin -> ghost copies."""
    dec = built.Decomp(dims, 8)
    dec.rank_map([1, 1, 1], [0, 0, 0])
    ro, so, nb, rp, sp = dec.exchange_plan(512)
    ex = built.Exchange(ctx, 0, ro, so, nb, rp, sp)
    field = built.uniform(dec.nbricks * 512, seed=77)
    d = ctx.upload(field)
    ex.run(d, built.STREAM_COMM)
    ctx.sync()
    got = ctx.download(d, dec.nbricks * 512)
    odec = oracle.Decomp(dims, 8)
    exp = oracle.exchange([field.copy()], odec, [1, 1, 1])[0]
    assert np.array_equal(got.view(np.uint64), exp.view(np.uint64))
    # only ghost bricks changed
    s1 = int(dec.sep_pos[1]) * 512
    assert np.array_equal(got[:s1], field[:s1])
    ex.close()
    ctx.free(d)


def _periodic_reference(oracle, which, field3d, sweeps):
    """`sweeps` applications of the stencil on a fully periodic array, via the oracle's array
    form on a halo-padded copy."""
    n2, n1, n0 = field3d.shape
    cur = field3d
    for _ in range(sweeps):
        pad = np.pad(cur, 8, mode="wrap")
        out = np.zeros_like(pad)
        oracle.array_sweep(which, pad, out, (8, 8 + n2, 8, 8 + n1, 8, 8 + n0))
        cur = out[8:-8, 8:-8, 8:-8].copy()
    return cur


@pytest.mark.parametrize("which", (7, 13, 25, 125))
@pytest.mark.parametrize("overlap", (False, True))
def test_weak_step_single_rank_equals_periodic_array(ctx, built, oracle, which, overlap):
    """exchange + ST_ITER sweeps on one periodic subdomain == ST_ITER periodic array sweeps: the
    end-to-end check of weak/main.cu:321-323 (brick result vs array result), independent of the
    brick layout, for both the serial and the overlapped schedule."""
    from bricklib_b200 import weak
    dims = (32, 48, 40)
    sub = weak.WeakSubdomain(ctx, dims, which)
    f = sub.host_field()
    sub.upload(f)
    sub.step(overlap=overlap)
    sub.step(overlap=overlap)
    ctx.sync()
    got = sub.download()
    dec = sub.dec
    t = dec.grid.shape
    ext = [t[2] * 8, t[1] * 8, t[0] * 8]
    arr0 = np.zeros((ext[2], ext[1], ext[0]))
    oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], arr0, dec.grid, f)
    exp = _periodic_reference(oracle, which, arr0[8:-8, 8:-8, 8:-8].copy(), 2 * sub.st_iter)
    arr1 = np.zeros_like(arr0)
    oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], arr1, dec.grid, got)
    assert relerr(arr1[8:-8, 8:-8, 8:-8], exp) <= REL_TOL
    sub.close()


def _ranks_in_threads(built, nranks, fn):
    """Run fn(rank, ctx) on `nranks` devices, one host thread per device, sharing a NCCL id."""
    uid = built.unique_id()
    errs, outs = [], [None] * nranks

    def work(r):
        try:
            c = built.Context(r)
            c.comm_init(r, nranks, uid)
            outs[r] = fn(r, c)
            c.sync()
            c.close()
        except Exception as e:  # noqa: BLE001
            errs.append((r, repr(e)))

    th = [threading.Thread(target=work, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    return outs


def _ndev(built):
    import ctypes
    n = ctypes.c_int()
    built.check(built.lib().bricks_device_count(ctypes.byref(n)))
    return n.value


@pytest.mark.parametrize("nranks", (2, 4, 8))
def test_multi_gpu_exchange_matches_oracle_bit_exact(built, oracle, nranks):
    if _ndev(built) < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    dims = (32, 32, 32)
    procdims = built.cart_dims(nranks)
    decs = [built.Decomp(dims, 8) for _ in range(nranks)]
    fields = [built.uniform(decs[0].nbricks * 512, seed=500 + r) for r in range(nranks)]

    def fn(r, c):
        dec = decs[r]
        dec.rank_map(procdims, built.cart_coords(r, procdims))
        ex = built.Exchange(c, r, *dec.exchange_plan(512))
        d = c.upload(fields[r])
        ex.run(d, built.STREAM_COMM)
        c.sync()
        out = c.download(d, dec.nbricks * 512)
        ex.close()
        c.free(d)
        return out

    got = _ranks_in_threads(built, nranks, fn)
    odec = oracle.Decomp(dims, 8)
    exp = oracle.exchange([f.copy() for f in fields], odec, procdims)
    for r in range(nranks):
        assert np.array_equal(got[r].view(np.uint64), exp[r].view(np.uint64)), r


@pytest.mark.parametrize("early", (False, True))
@pytest.mark.parametrize("nranks", (2, 8))
def test_multi_gpu_weak_step_equals_global_periodic_array(built, oracle, nranks, early):
    """Weak scaling end to end: P subdomains on a periodic process grid, exchange over NCCL,
    ST_ITER sweeps with the overlapped schedule == periodic sweeps of the assembled global array."""
    if _ndev(built) < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    from bricklib_b200 import weak
    dims, which = (32, 32, 32), 7
    procdims = built.cart_dims(nranks)
    subs = [None] * nranks
    f0 = [None] * nranks

    def fn(r, c):
        sub = weak.WeakSubdomain(c, dims, which, rank=r, procdims=procdims)
        subs[r] = sub
        f0[r] = sub.host_field()
        sub.upload(f0[r])
        for _ in range(nsteps):
            sub.step(overlap=True, early=early)
        c.sync()
        out = sub.download()
        return out

    nsteps = 2 if early else 1
    got = _ranks_in_threads(built, nranks, fn)
    odec = oracle.Decomp(dims, 8)
    ext = [48, 48, 48]

    def interior_of(field):
        a = np.zeros((48, 48, 48))
        oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], a, odec.grid, field)
        return a[8:-8, 8:-8, 8:-8]

    # process grid index 0 is the slowest (k) axis: coo = (ck, cj, ci); populate() puts the "+d"
    # neighbour at coordinate - 1 (brick-mpi.h:739-751), so physical position runs against coo
    pk, pj, pi = procdims

    def where(r):
        c = built.cart_coords(r, procdims)
        return [(-c[a]) % procdims[a] for a in range(3)]

    glob = np.zeros((32 * pk, 32 * pj, 32 * pi))
    for r in range(nranks):
        ck, cj, ci = where(r)
        glob[32 * ck:32 * ck + 32, 32 * cj:32 * cj + 32, 32 * ci:32 * ci + 32] = interior_of(f0[r])
    exp = _periodic_reference(oracle, which, glob, nsteps * built.ST_ITER[which])
    for r in range(nranks):
        ck, cj, ci = where(r)
        e = exp[32 * ck:32 * ck + 32, 32 * cj:32 * cj + 32, 32 * ci:32 * ci + 32]
        assert relerr(interior_of(got[r]), e) <= REL_TOL, r


@pytest.mark.parametrize("kind", ("regions", "packed", "shift"))
@pytest.mark.parametrize("nranks", (1, 2, 4, 8))
def test_exchange_kinds_fill_the_ghost_zone_like_the_oracle(built, oracle, nranks, kind):
    """every form of the exchange the step object can run -- one send/recv pair per region
    (BrickDecomp::exchange, brick-mpi.h:466-495), one packed message per peer (the role of
    ExchangeView, brick-mpi.h:96-123) and the SHIFT exchange (brick-mpi.h:587-681) -- leaves the
    storage bit-identical to the oracle's emulation on `nranks` virtual ranks"""
    if _ndev(built) < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    from bricklib_b200 import weak
    dims = (32, 40, 48)
    procdims = built.cart_dims(nranks)
    nb = built.Decomp(dims, 8).nbricks
    fields = [built.uniform(nb * 512, seed=700 + r) for r in range(nranks)]

    def fn(r, c):
        sub = weak.WeakSubdomain(c, dims, 7, rank=r, procdims=procdims, exchange=kind)
        sub.upload(fields[r])
        sub.exchange_only()
        c.sync()
        out = sub.download()
        sub.close()
        return out

    got = _ranks_in_threads(built, nranks, fn)
    odec = oracle.Decomp(dims, 8)
    exp = oracle.exchange([f.copy() for f in fields], odec, procdims)
    for r in range(nranks):
        assert np.array_equal(got[r].view(np.uint64), exp[r].view(np.uint64)), r


def test_plan_rejects_an_adjacency_that_contradicts_the_grid(ctx, built):
    """the marching kernels take the +i / +-k neighbours from the grid, the reference accessor from
    adj (brick.h:229-246): a BrickInfo that is not lattice-consistent is an error, not a silent
    difference"""
    dec = built.Decomp((16, 16, 16), 8)
    adj = dec.adj.copy()
    b = int(dec.grid[1, 1, 1])
    adj[b, 14] = int(dec.grid[2, 2, 2])
    with pytest.raises(built.BricksError, match="consistent with the grid"):
        built.Plan(ctx, dec.grid, adj)
    dec.close()


def test_derived_plans_share_tables_and_give_the_same_sweep(ctx, built):
    """bricks_plan_derive: a plan over other boxes of the parent's grid borrows its device tables"""
    dec = built.Decomp((32, 32, 32), 8)
    g = tuple(int(x) for x in dec.gdims)
    parent = built.Plan(ctx, dec.grid, dec.adj)
    own = [((1, 1, 1), tuple(x - 1 for x in g))]
    child = built.Plan(ctx, boxes=own, parent=parent)
    alone = built.Plan(ctx, dec.grid, dec.adj, own)
    assert child.device_tables() == parent.device_tables() != alone.device_tables()
    f = built.uniform(dec.nbricks * 512, seed=4)
    d_in = ctx.upload(f)
    outs = []
    for p in (child, alone):
        d = ctx.upload(np.zeros_like(f))
        p.stencil(125, d_in, d)
        ctx.sync()
        outs.append(ctx.download(d, dec.nbricks * 512))
        ctx.free(d)
    assert np.array_equal(outs[0].view(np.uint64), outs[1].view(np.uint64)) and outs[0].any()
    child.close()
    alone.close()
    parent.close()
    ctx.free(d_in)
    dec.close()


def test_host_steps_pipeline_matches_plain_step(ctx, built):
    """the software-pipelined host-fed steps (two device pairs, copy streams) give the same field
    as upload -> step -> download, for every step of the pipeline"""
    from bricklib_b200 import weak
    sub = weak.WeakSubdomain(ctx, (32, 32, 32), 125)
    f = sub.host_field()
    sub.upload(f)
    sub.step(overlap=True)
    ctx.sync()
    want = sub.download()
    outs = [np.zeros_like(f), np.zeros_like(f)]
    sub.host_steps(f, outs, 5, overlap=True)
    ctx.sync()
    n = int(sub.dec.sep_pos[1]) * 512  # owned bricks; ghost bricks never cross the bus
    assert sub.host_step_bytes == n * 8
    for o in outs:
        assert np.array_equal(o[:n].view(np.uint64), want[:n].view(np.uint64))
        assert not o[n:].any()
    sub.close()


@pytest.mark.parametrize("which", (7, 25, 125))
@pytest.mark.parametrize("kind", ("regions", "packed"))
def test_early_exchange_schedule_is_bitwise_the_plain_schedule(ctx, built, which, kind):
    """starting the next step's exchange right after the last sweep's skin bricks (early), with the
    following first sweep split or unsplit, changes the order of launches only: three steps give
    the same bits on the owned bricks as the serial schedule"""
    from bricklib_b200 import weak
    sub = weak.WeakSubdomain(ctx, (32, 40, 48), which, exchange=kind)
    f = sub.host_field()
    res = []
    for overlap, early, unsplit in ((False, False, False), (True, False, False), (True, True, False), (True, True, True)):
        sub.upload(f)
        for _ in range(3):
            sub.step(overlap=overlap, early=early, unsplit=unsplit)
        ctx.sync()
        res.append(sub.download())
    n = int(sub.dec.sep_pos[1]) * 512
    for r in res[1:]:
        assert np.array_equal(res[0][:n].view(np.uint64), r[:n].view(np.uint64))
    sub.close()


def test_step_trace_gives_one_record_per_operation(ctx, built):
    """bricks_step_trace: the timeline of a step (the per-phase timers of weak/main.cu:251-286):
    a self exchange, the split first sweep, the middle sweeps and the owned-bricks last sweep"""
    from bricklib_b200 import weak
    sub = weak.WeakSubdomain(ctx, (64, 64, 64), 125)
    sub.upload(sub.host_field())
    sub.step_obj.trace(True)
    sub.step(overlap=True)
    tl = sub.step_obj.trace_read()
    sub.step_obj.trace(False)
    names = [n for n, _, _ in tl]
    # the sweep an exchange hides behind is enqueued before the exchange's (many) host calls
    assert names == ["sweep 1 inner", "comm[0] exchange", "sweep 1 shell", "sweep 2 full", "sweep 3 full", "sweep 4 owned"]
    assert all(0.0 <= a <= b for _, a, b in tl) and min(a for _, a, _ in tl) == 0.0
    sub.step(overlap=True)  # untraced steps leave the last timeline alone
    assert [n for n, _, _ in sub.step_obj.trace_read()] == names
    sub.close()


def _thread_allgather(nranks):
    """allgather for ranks that are threads of this process: a slot per rank and a barrier"""
    slots, bar = [None] * nranks, threading.Barrier(nranks)

    def make(r):
        def allgather(b):
            slots[r] = b
            bar.wait()
            out = list(slots)
            bar.wait()
            return out
        return allgather
    return make, bar


@pytest.mark.parametrize("ce_min", (None, 16384))
@pytest.mark.parametrize("kind", ("regions", "shift"))
@pytest.mark.parametrize("nranks", (2, 4, 8))
def test_copy_engine_exchange_is_bit_exact_and_steps_like_the_global_array(built, oracle, monkeypatch, nranks, kind, ce_min):
    """bricks_step_p2p_*: the region messages as cudaMemcpyAsync into the peers' ghost bricks, ordered
    by flag words.  (1) one bare exchange leaves every rank's storage bit-identical to the oracle's
    emulation of BrickDecomp::exchange; (2) three steps with the early schedule == periodic sweeps of
    the assembled global array (weak/main.cu:321-323)."""
    if _ndev(built) < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    if ce_min is not None:
        # regions of >= 16 KiB (the faces of a 32^3 subdomain) by copy engine, edges and corners in the
        # store kernel: both halves of the production path (threshold 1 MiB) in one exchange
        monkeypatch.setenv("BRICKS_P2P_CE_MIN", str(ce_min))
    else:
        monkeypatch.delenv("BRICKS_P2P_CE_MIN", raising=False)
    from bricklib_b200 import weak
    dims, which, nsteps = (32, 32, 32), 13, 3
    procdims = built.cart_dims(nranks)
    nb = built.Decomp(dims, 8).nbricks
    fields = [built.uniform(nb * 512, seed=900 + r) for r in range(nranks)]
    make, bar = _thread_allgather(nranks)
    f0 = [None] * nranks
    stepped = [None] * nranks

    def fn(r, c):
        sub = weak.WeakSubdomain(c, dims, which, rank=r, procdims=procdims, exchange=kind)
        sub.enable_p2p(make(r))
        assert "peer-mapped ghost bricks" in sub.exchange_kind
        sub.upload(fields[r])
        bar.wait()
        sub.exchange_only()
        c.sync()
        out = sub.download()
        bar.wait()
        f0[r] = sub.host_field()
        sub.upload(f0[r])
        bar.wait()
        for _ in range(nsteps):
            sub.step(overlap=True, early=True)
        c.sync()
        stepped[r] = sub.download()
        bar.wait()  # nobody unmaps while a peer may still copy
        sub.close()
        return out

    got = _ranks_in_threads(built, nranks, fn)
    odec = oracle.Decomp(dims, 8)
    exp = oracle.exchange([f.copy() for f in fields], odec, procdims)
    for r in range(nranks):
        assert np.array_equal(got[r].view(np.uint64), exp[r].view(np.uint64)), r

    ext = [48, 48, 48]

    def interior_of(field):
        a = np.zeros((48, 48, 48))
        oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], a, odec.grid, field)
        return a[8:-8, 8:-8, 8:-8]

    pk, pj, pi = procdims

    def where(r):
        c = built.cart_coords(r, procdims)
        return [(-c[a]) % procdims[a] for a in range(3)]

    glob = np.zeros((32 * pk, 32 * pj, 32 * pi))
    for r in range(nranks):
        ck, cj, ci = where(r)
        glob[32 * ck:32 * ck + 32, 32 * cj:32 * cj + 32, 32 * ci:32 * ci + 32] = interior_of(f0[r])
    want = _periodic_reference(oracle, which, glob, nsteps * built.ST_ITER[which])
    for r in range(nranks):
        ck, cj, ci = where(r)
        e = want[32 * ck:32 * ck + 32, 32 * cj:32 * cj + 32, 32 * ci:32 * ci + 32]
        assert relerr(interior_of(stepped[r]), e) <= REL_TOL, r
