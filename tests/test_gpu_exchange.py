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
def test_early_exchange_schedule_is_bitwise_the_plain_schedule(ctx, built, which):
    """starting the next step's exchange right after the last sweep's skin bricks (early=True)
    changes the order of launches only: three steps give the same bits on the owned bricks"""
    from bricklib_b200 import weak
    sub = weak.WeakSubdomain(ctx, (32, 40, 48), which)
    f = sub.host_field()
    res = []
    for early in (False, True):
        sub.upload(f)
        for _ in range(3):
            sub.step(overlap=True, early=early)
        ctx.sync()
        res.append(sub.download())
    n = int(sub.dec.sep_pos[1]) * 512
    assert np.array_equal(res[0][:n].view(np.uint64), res[1][:n].view(np.uint64))
    sub.close()
