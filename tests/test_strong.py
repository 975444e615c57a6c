"""Strong-scaling (two-level decomposition) tests.  CPU part: the product's C++ planner
(include/bricks/strong-plan.h) against the oracle's restatement of strong/args.cpp:36-56,104-113
and strong/main.cu:188-247.  GPU part: exchange bit-exact, full step vs a periodic array."""
import threading

import numpy as np
import pytest

from conftest import REL_TOL, relerr


@pytest.mark.parametrize("nranks", (1, 2, 3, 8))
def test_planner_matches_oracle_restatement(built, oracle, nranks):
    dec = built.Decomp((32, 32, 32), 8)
    odec = oracle.Decomp((32, 32, 32), 8)
    subdim, total = 4, 64
    all_recv = {}
    for rk in range(nranks):
        p = built.StrongPlan(128, 32, nranks, rk, dec)
        assert (p.first_id, p.last_id) == oracle.strong_range(rk, total, nranks)
        # every (receiver subdomain, region) appears exactly once: as a local link or as a received piece
        seen = set()
        for ss, ds, so, do, n in p.links:
            rid = [i for i in range(len(odec.ghost)) if int(odec.ghost[i]["pos"]) * 4096 == do]
            assert len(rid) == 1
            i = rid[0]
            sender = p.first_id + ss
            assert sender == oracle.strong_neighbor(p.first_id + ds, int(odec.ghost[i]["neighbor"]), subdim)
            assert so == int(odec.skin[i]["pos"]) * 4096 and n == int(odec.ghost[i]["len"]) * 4096
            seen.add((p.first_id + ds, i))
        for peer in p.recv:
            pos = 0
            for sub, region, off, n, buf_off, sender in peer["pieces"]:
                assert buf_off == pos
                pos += n
                assert oracle.strong_owner(sender, total, nranks) == peer["rank"]
                assert sender == oracle.strong_neighbor(p.first_id + sub, int(odec.ghost[region]["neighbor"]), subdim)
                assert off == int(odec.ghost[region]["pos"]) * 4096
                seen.add((p.first_id + sub, region))
                all_recv[(rk, peer["rank"])] = [(s, r, n) for _, r, _, n, _, s in peer["pieces"]]
            assert pos == peer["bytes"]
        assert len(seen) == p.nsub * len(odec.ghost)
        p.close()
    # sender and receiver enumerate the pieces of a message in the same order
    for rk in range(nranks):
        p = built.StrongPlan(128, 32, nranks, rk, dec)
        for peer in p.send:
            mine = [(s, r, n) for _, r, _, n, _, s in peer["pieces"]]
            assert mine == all_recv[(peer["rank"], rk)]
        p.close()
    dec.close()


def test_planner_rejects_bad_shapes(built):
    dec = built.Decomp((32, 32, 32), 8)
    with pytest.raises(built.BricksError):
        built.StrongPlan(96, 32, 1, 0, dec)  # 3 subdomains per axis: not a power of two
    with pytest.raises(built.BricksError):
        built.StrongPlan(64, 32, 9, 0, dec)  # more ranks than subdomains
    dec.close()


def _global_from(oracle, odec, fields_by_id, subdim, sdom):
    """assemble the interiors of all subdomains into the global periodic array [k][j][i]"""
    ext = [sdom + 16] * 3
    glob = np.zeros((subdim * sdom,) * 3)
    for sid, f in fields_by_id.items():
        a = np.zeros((sdom + 16,) * 3)
        oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], a, odec.grid, f)
        ci, cj, ck = oracle.morton3_coords(sid)
        glob[ck * sdom:(ck + 1) * sdom, cj * sdom:(cj + 1) * sdom, ci * sdom:(ci + 1) * sdom] = a[8:-8, 8:-8, 8:-8]
    return glob


def _periodic_reference(oracle, which, field3d, sweeps):
    cur = field3d
    n2, n1, n0 = cur.shape
    for _ in range(sweeps):
        pad = np.pad(cur, 8, mode="wrap")
        out = np.zeros_like(pad)
        oracle.array_sweep(which, pad, out, (8, 8 + n2, 8, 8 + n1, 8, 8 + n0))
        cur = out[8:-8, 8:-8, 8:-8].copy()
    return cur


@pytest.mark.gpu
@pytest.mark.parametrize("dom,sdom", [(64, 32), (64, 16)])
def test_strong_exchange_single_gpu_bit_exact(ctx, built, oracle, dom, sdom):
    from bricklib_b200 import strong
    sd = strong.StrongDomain(ctx, dom, sdom, 7)
    fields = {sd.plan.first_id + s: sd.host_field(s) for s in range(sd.nsub)}
    sd.upload()
    sd.exchange_only()
    ctx.sync()
    odec = oracle.Decomp((sdom,) * 3, 8)
    exp = oracle.strong_exchange({k: v.copy() for k, v in fields.items()}, odec, dom // sdom)
    for s in range(sd.nsub):
        got = sd.download(s)
        assert np.array_equal(got.view(np.uint64), exp[sd.plan.first_id + s].view(np.uint64)), s
    sd.close()


@pytest.mark.gpu
@pytest.mark.parametrize("which", (7, 125))
@pytest.mark.parametrize("overlap", (False, True))
def test_strong_step_single_gpu_equals_periodic_array(ctx, built, oracle, which, overlap):
    from bricklib_b200 import strong
    dom, sdom = 64, 32
    sd = strong.StrongDomain(ctx, dom, sdom, which)
    fields = {sd.plan.first_id + s: sd.host_field(s) for s in range(sd.nsub)}
    sd.upload()
    n = sd.step(overlap=overlap)
    assert n >= sd.st_iter + 1
    ctx.sync()
    odec = oracle.Decomp((sdom,) * 3, 8)
    exp = _periodic_reference(oracle, which, _global_from(oracle, odec, fields, dom // sdom, sdom), sd.st_iter)
    got = _global_from(oracle, odec, {sd.plan.first_id + s: sd.download(s) for s in range(sd.nsub)}, dom // sdom, sdom)
    assert relerr(got, exp) <= REL_TOL
    sd.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", (2, 8))
def test_strong_step_multi_gpu_equals_periodic_array(built, oracle, nranks):
    import ctypes
    n = ctypes.c_int()
    built.check(built.lib().bricks_device_count(ctypes.byref(n)))
    if n.value < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    from bricklib_b200 import strong
    dom, sdom, which = 128, 32, 125
    uid = built.unique_id()
    fields, results, errs = {}, {}, []

    def work(r):
        try:
            c = built.Context(r)
            c.comm_init(r, nranks, uid)
            sd = strong.StrongDomain(c, dom, sdom, which, rank=r, nranks=nranks)
            for s in range(sd.nsub):
                fields[sd.plan.first_id + s] = sd.host_field(s)
            sd.upload()
            sd.step(overlap=True)
            c.sync()
            for s in range(sd.nsub):
                results[sd.plan.first_id + s] = sd.download(s)
            c.sync()
        except Exception as e:  # noqa: BLE001
            errs.append((r, repr(e)))

    th = [threading.Thread(target=work, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    odec = oracle.Decomp((sdom,) * 3, 8)
    exp = _periodic_reference(oracle, which, _global_from(oracle, odec, fields, dom // sdom, sdom), built.ST_ITER[which])
    got = _global_from(oracle, odec, results, dom // sdom, sdom)
    assert relerr(got, exp) <= REL_TOL
