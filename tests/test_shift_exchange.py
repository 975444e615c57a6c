"""The SHIFT (multi-stage) exchange -- BrickDecomp::multiStageExchangeView, include/brick-mpi.h:587-681:
one stage per axis, corner and edge data forwarded through the face neighbours' ghost zones.
CPU: (1) the oracle's restatement fills every ghost brick exactly as the direct exchange does,
(2) the product's piece tables (BrickDecomp::shiftPieces through bricks_decomp_shift_plan) move the
same bytes as the oracle's packages, on 1..8 virtual ranks.  GPU: the three stages run through
bricks_exchange_* and must leave a storage bit-identical to the direct exchange."""
import threading

import numpy as np
import pytest

GRIDS = [[1, 1, 1], [2, 1, 1], [2, 2, 1], [2, 2, 2], [3, 2, 1], [4, 1, 1]]
DIMS = [(16, 16, 16), (32, 24, 16), (24, 40, 32)]


def _fields(nelem, nranks, seed):
    rng = np.random.default_rng(seed)
    return [rng.random(nelem) for _ in range(nranks)]


@pytest.mark.parametrize("dims", DIMS)
def test_oracle_shift_exchange_equals_direct_exchange(oracle, dims):
    dec = oracle.Decomp(dims, 8)
    stages = oracle.multistage_packages(dec)
    assert [len(p) for p in stages] == [2, 2, 2]  # two messages per stage, six in all
    nghost = dec.nbricks - int(dec.sep_pos[1])
    assert sum(l for p in stages for _, recv, _ in p for _, l in recv) == nghost
    for procdims in GRIDS:
        f = _fields(dec.nbricks * 512, int(np.prod(procdims)), 3)
        direct = oracle.exchange([x.copy() for x in f], dec, procdims)
        shifted = oracle.multistage_exchange([x.copy() for x in f], dec, procdims)
        for a, b in zip(direct, shifted):
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def _run_plan_on_virtual_ranks(built, dims, procdims, fields):
    """Apply the product's stage tables the way NCCL would: per stage, message i of rank r is
    received from recv_peer[i]; that peer's i-th piece towards r (pieces to one peer are matched in
    issue order) is the source."""
    nranks = len(fields)
    plans = []
    for r in range(nranks):
        dec = built.Decomp(dims, 8)
        dec.rank_map(procdims, built.cart_coords(r, procdims))
        plans.append(dec.shift_plan(512))
        dec.close()
    for stage in range(3):
        snap = [f.view(np.uint8).copy() for f in fields]
        for r in range(nranks):
            ro, so, nb, rp, sp = plans[r][stage]
            taken = {}
            for i in range(len(ro)):
                src = int(rp[i])
                pro, pso, pnb, prp, psp = plans[src][stage]
                k = taken.get(src, 0)
                sends = [j for j in range(len(pso)) if int(psp[j]) == r]
                # my i-th receive from `src` pairs with src's k-th send to me
                j = sends[k]
                taken[src] = k + 1
                assert int(pnb[j]) == int(nb[i])
                dst = fields[r].view(np.uint8)
                dst[int(ro[i]):int(ro[i]) + int(nb[i])] = snap[src][int(pso[j]):int(pso[j]) + int(pnb[j])]
    return fields


@pytest.mark.parametrize("dims", DIMS)
def test_product_shift_plan_moves_the_same_bytes_as_the_oracle(built, oracle, dims):
    odec = oracle.Decomp(dims, 8)
    for procdims in GRIDS:
        f = _fields(odec.nbricks * 512, int(np.prod(procdims)), 5)
        exp = oracle.multistage_exchange([x.copy() for x in f], odec, procdims)
        got = _run_plan_on_virtual_ranks(built, dims, procdims, [x.copy() for x in f])
        for a, b in zip(exp, got):
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def test_product_shift_plan_is_six_messages_of_contiguous_pieces(built, oracle):
    dec = built.Decomp((512, 512, 512), 8)
    dec.rank_map([2, 2, 2], [0, 1, 0])
    stages = dec.shift_plan(512)
    direct = dec.exchange_plan(512)
    total = 0
    for ro, so, nb, rp, sp in stages:
        assert len(set(zip(rp.tolist(), sp.tolist()))) <= 2  # (from +d, to -d) and (from -d, to +d)
        assert np.all(nb > 0) and np.all(ro % 4096 == 0) and np.all(so % 4096 == 0)
        total += int(nb.sum())
    assert total == int(direct[2].sum())  # every ghost byte arrives exactly once
    # receive ranges tile the ghost zone without overlap
    spans = sorted((int(a), int(a) + int(n)) for ro, so, nb, rp, sp in stages for a, n in zip(ro, nb))
    assert all(spans[i][1] <= spans[i + 1][0] for i in range(len(spans) - 1))
    assert spans[0][0] == int(dec.sep_pos[1]) * 4096 and spans[-1][1] == dec.nbricks * 4096
    dec.close()


# ------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(16, 16, 16), (32, 24, 16), (64, 64, 64), (512, 512, 512)])
def test_gpu_self_shift_exchange_equals_direct_exchange(ctx, built, dims):
    """np = 1 on a periodic grid: all six messages are device copies, stage after stage."""
    dec = built.Decomp(dims, 8)
    dec.rank_map([1, 1, 1], [0, 0, 0])
    direct = built.Exchange(ctx, 0, *dec.exchange_plan(512))
    shift = built.ShiftExchange(ctx, 0, dec.shift_plan(512))
    field = built.uniform(dec.nbricks * 512, seed=21)
    d1, d2 = ctx.upload(field), ctx.upload(field)
    direct.run(d1, built.STREAM_COMM)
    shift.run(d2, built.STREAM_COMM)
    ctx.sync()
    a, b = ctx.download(d1, dec.nbricks * 512), ctx.download(d2, dec.nbricks * 512)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    s1 = int(dec.sep_pos[1]) * 512
    assert np.array_equal(b[:s1], field[:s1])  # only ghost bricks change
    for p in (d1, d2):
        ctx.free(p)
    direct.close()
    shift.close()
    dec.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", (2, 4, 8))
def test_multi_gpu_shift_exchange_matches_oracle_bit_exact(built, oracle, nranks):
    """One rank per GPU over NCCL: after the three stages every rank's storage equals the oracle's
    direct exchange (and so its SHIFT restatement) bit for bit."""
    from test_gpu_exchange import _ndev, _ranks_in_threads
    if _ndev(built) < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    dims = (32, 32, 32)
    procdims = built.cart_dims(nranks)
    decs = [built.Decomp(dims, 8) for _ in range(nranks)]
    fields = [built.uniform(decs[0].nbricks * 512, seed=700 + r) for r in range(nranks)]

    def fn(r, c):
        dec = decs[r]
        dec.rank_map(procdims, built.cart_coords(r, procdims))
        ex = built.ShiftExchange(c, r, dec.shift_plan(512))
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
