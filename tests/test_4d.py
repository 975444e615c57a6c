"""The 4-D path (weak/main-4d.cpp in the reference): BrickDecomp<4,4,4,4,4>
numbering, the 9-point star of stencils/mpi9pt.py on 4x4x4x4 bricks, exchange + 8 shrinking sweeps.

CPU part: the oracle (oracle/oracle4d.py, oracle/stencil4d_oracle.c) against the golden fixtures dumped
from the compiled reference and against oracle/_ref itself when present; the product's own numbering
(include/bricks/brick-decomp.h through bricks_decomp4d_*) against both; the kernel bodies of
bricklib_b200/csrc/stencil4d_body.cuh emulated thread by thread on the host against the oracle.
GPU part: the kernels through the C ABI against the oracle, in a worker process (tests/gpu4d_worker.py).
"""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, REL_TOL, ROOT, fnv1a64, relerr

D4_STENCIL_DIMS = [(16, 16, 16, 16), (16, 24, 16, 20)]


@pytest.fixture(scope="module")
def o4(oracle):
    from oracle import oracle4d
    return oracle4d


@pytest.fixture(scope="module")
def golden4():
    with open(os.path.join(GOLDEN, "decomp4d.json")) as f:
        return json.load(f)["decomp"]


def _check(dec_regions, nbricks, sep_pos, skin_size, grid, adj, rec):
    assert nbricks == rec["nbricks"]
    assert [int(x) for x in sep_pos] == rec["sep_pos"]
    assert [int(x) for x in skin_size] == rec["skin_size"]
    for name in ("ghost", "skin"):
        for k, v in rec[name].items():
            assert [int(x) for x in dec_regions(name)[k]] == v, (name, k)
    if nbricks <= 7000:  # hashing in pure Python is slow; the larger cases go through tables above
        assert fnv1a64(grid) == rec["grid_fnv"]
        assert fnv1a64(adj) == rec["adj_fnv"]


@pytest.mark.parametrize("idx", range(10))
def test_oracle_numbering_matches_reference_fixture(o4, golden4, idx):
    rec = golden4[idx]
    dec = o4.DecompND(rec["dims"], depth=rec["gz"], page_align=rec["page_align"])
    _check(dec.regions, dec.nbricks, dec.sep_pos, dec.skin_size, dec.grid, dec.adj, rec)


@pytest.mark.parametrize("idx", range(10))
def test_product_numbering_matches_reference_fixture(built, golden4, idx):
    """BrickDecomp<4,...> of include/bricks/brick-decomp.h: page-aligned and dense numbering, ghost depth of
    one and two bricks (padding only appears with odd region sizes)"""
    rec = golden4[idx]
    dec = built.Decomp4D(rec["dims"], rec["gz"], rec["page_align"])
    regions = lambda name: {k: (dec.ghost if name == "ghost" else dec.skin)[k] for k in rec[name]}  # noqa: E731
    _check(regions, dec.nbricks, dec.sep_pos, dec.skin_size, dec.grid, dec.adj, rec)
    dec.close()


def test_rank_maps_match_reference_fixture(built, o4, golden4):
    rec = golden4[0]
    dec = built.Decomp4D(rec["dims"], rec["gz"], True)
    for case in rec["rank_map"]:
        want = {int(k, 16): v for k, v in case["map"].items()}
        assert o4.rank_map(case["procdims"], case["coo"]) == want
        assert dec.rank_map(case["procdims"], case["coo"]) == want
    dec.close()


def test_numbering_against_compiled_reference_when_present(built, o4):
    if not o4.ref_available():
        pytest.skip("oracle/_ref/libref4d*.so not built on this machine")
    for dims, gz in [((32, 24, 20, 28), 8), ((20, 12, 16, 24), 4)]:
        for pa in (True, False):
            ref = o4.Ref4D(dims, gz=gz, page_align=pa)
            mine = built.Decomp4D(dims, gz, pa)
            orc = o4.DecompND(dims, depth=gz, page_align=pa)
            for got in (mine, orc):
                assert got.nbricks == ref.nbricks
                assert np.array_equal(got.grid, ref.grid) and np.array_equal(got.adj, ref.adj)
            for name, tab in (("ghost", mine.ghost), ("skin", mine.skin)):
                for k, v in ref.regions[name].items():
                    assert np.array_equal(tab[k].astype(np.uint64), v.astype(np.uint64)), (name, k)
            ref.close()
            mine.close()


@pytest.mark.parametrize("dims", D4_STENCIL_DIMS)
def test_oracle_stencil_matches_reference_outputs(o4, dims):
    """the generated kernel of the reference (AVX2, fold 2,2) on the three sweep regions of a period"""
    from bricklib_b200 import synth
    fx = np.load(os.path.join(GOLDEN, "d4pt9_%dx%dx%dx%d.npz" % dims))
    dec = o4.DecompND(dims, page_align=False)
    inp = synth.uniform(dec.nbricks * 256, seed=17)
    for s in (0, 1, 2):
        blist = o4.sweep_list(dec.grid, s)
        assert len(blist) == int(fx["n_%d" % s][0])
        out = np.zeros(dec.nbricks * 256)
        o4.brick_sweep(dec.adj, blist, inp, out)
        pick = fx["bricks_%d" % s]
        assert set(pick.tolist()) <= set(blist.tolist())
        assert relerr(out.reshape(-1, 256)[pick], fx["brick_%d" % s]) <= 1e-14
        if o4.ref_available():  # and the whole field against the reference itself
            ref = np.zeros_like(out)
            o4.ref_brick_sweep(dec.adj, blist, inp, ref)
            assert relerr(out, np.where(ref == 0, out, ref)) <= 1e-14 and np.array_equal(out == 0, ref == 0)


def test_oracle_brick_and_array_forms_agree(o4):
    """the reference's own self-check (compareBrick of the array result against the brick result,
    weak/main-4d.cpp:347-350) at the strict tolerance"""
    from bricklib_b200 import synth
    dec = o4.DecompND((16, 24, 16, 20), page_align=False)
    n = dec.nbricks * 256
    inp = synth.uniform(n, seed=5)
    t = [x * 4 for x in dec.t]
    arr = np.zeros([x + 8 for x in t[::-1]])
    o4.copy_brick(False, t, (4,) * 4, (0,) * 4, arr, dec.grid, inp)
    aout = np.zeros_like(arr)
    o4.array_sweep(arr, aout, [8] * 4, [x for x in t])
    bout = np.zeros(n)
    o4.brick_sweep(dec.adj, o4.sweep_list(dec.grid, 1), inp, bout)
    back = np.zeros_like(arr)
    o4.copy_brick(False, t, (4,) * 4, (0,) * 4, back, dec.grid, bout)
    box = tuple(slice(8, x) for x in t[::-1])
    assert relerr(back[box], aout[box]) <= REL_TOL and aout[box].any()


def test_product_copy_brick4_and_exchange_plan(built, o4):
    dims = (16, 24, 16, 20)
    dec = built.Decomp4D(dims)
    odec = o4.DecompND(dims, page_align=False)
    t = [g * 4 for g in dec.gdims]
    rng = np.random.default_rng(3)
    arr = rng.random([x + 8 for x in t[::-1]])
    b1, b2 = np.zeros(dec.nbricks * 256), np.zeros(dec.nbricks * 256)
    built.copy_brick4(True, t, (4,) * 4, (0,) * 4, arr, dec.grid, b1, dec.adj)
    o4.copy_brick(True, t, (4,) * 4, (0,) * 4, arr, dec.grid, b2)
    assert np.array_equal(b1, b2) and b1.any()
    back = np.zeros_like(arr)
    built.copy_brick4(False, list(dims), (4,) * 4, (8,) * 4, back, dec.grid, b1, dec.adj)
    inner = tuple(slice(12, 12 + x) for x in dims[::-1])
    assert np.array_equal(back[inner], arr[inner]) and not back[0].any()
    # the exchange plan is the oracle's region table and rank map in bytes
    procdims, coo = (2, 1, 2, 1), (1, 0, 1, 0)
    dec.rank_map(procdims, coo)
    ro, so, nb, rp, sp = dec.exchange_plan(256)
    rm = o4.rank_map(procdims, coo)
    for i, (g, s) in enumerate(zip(odec.ghost, odec.skin)):
        assert ro[i] == (g["pos"] + g["first_pad"]) * 2048 and so[i] == (s["pos"] + s["first_pad"]) * 2048
        assert nb[i] == (g["len"] - g["first_pad"] - g["last_pad"]) * 2048
        assert rp[i] == rm[g["neighbor"]] and sp[i] == rm[s["neighbor"]]
    dec.close()


def test_oracle_exchange_assembles_the_periodic_global_field(o4):
    """four virtual ranks on a (2,1,2,1) process grid: after the exchange every rank's storage, ghost bricks
    included, equals its window of the periodic GLOBAL array -- the sign convention of populate()
    (include/brick-mpi.h:730-753: "+d" is the neighbour at coordinate - 1) and the pairing of skin with ghost
    regions, checked against something that is not a region table"""
    dims, procdims, gz = (16, 16, 16, 16), (2, 1, 2, 1), 8
    P = int(np.prod(procdims))
    dec = o4.DecompND(dims, page_align=False)
    # array axis a (0 = contiguous) belongs to direction-set axis a + 1, i.e. to Cartesian coordinate coo[3 - a]
    pa = [procdims[3 - a] for a in range(4)]
    glob = np.random.default_rng(8).random([dims[a] * pa[a] for a in range(4)][::-1])
    ext = [d + 2 * gz for d in dims]
    stores, where = [], []
    for r in range(P):
        coo, rr = [], r
        for n in reversed(procdims):
            coo.append(rr % n)
            rr //= n
        coo = coo[::-1]
        blk = [(-coo[3 - a]) % pa[a] for a in range(4)]  # physical block of this rank along array axis a
        where.append(blk)
        own = glob[tuple(slice(blk[a] * dims[a], (blk[a] + 1) * dims[a]) for a in (3, 2, 1, 0))]
        arr = np.zeros(ext[::-1])
        arr[tuple(slice(gz, gz + dims[a]) for a in (3, 2, 1, 0))] = own
        st = np.zeros(dec.nbricks * 256)
        o4.copy_brick(True, dims, (0,) * 4, (gz,) * 4, arr, dec.grid, st)
        stores.append(st)
    o4.exchange(stores, dec, procdims)
    for r in range(P):
        back = np.zeros(ext[::-1])
        o4.copy_brick(False, ext, (0,) * 4, (0,) * 4, back, dec.grid, stores[r])
        idx = [(np.arange(-gz, dims[a] + gz) + where[r][a] * dims[a]) % (dims[a] * pa[a]) for a in range(4)]
        want = glob[np.ix_(idx[3], idx[2], idx[1], idx[0])]
        assert np.array_equal(back, want), r


def test_plan4d_rejects_bad_arguments(built):
    with pytest.raises(built.BricksError):
        built.Decomp4D((16, 16, 18, 16))  # not a multiple of the brick edge
    with pytest.raises(built.BricksError):
        built.Decomp4D((16, 16, 8, 16))   # smaller than two ghost depths


# ---------------------------------------------------------------- kernel bodies emulated on the host ----
@pytest.fixture(scope="module")
def emul(built):
    """tests/emul/stencil4d_emul.cpp: the kernel bodies compiled as host code (no FMA contraction, so the
    result is bit-comparable with the oracle)"""
    out = os.path.join(ROOT, "tests", "emul", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "stencil4d_emul.so")
    src = os.path.join(ROOT, "tests", "emul", "stencil4d_emul.cpp")
    hdr = os.path.join(ROOT, "bricklib_b200", "csrc", "stencil4d_body.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", so, src])
    L = C.CDLL(so)
    L.emul_stencil4d.argtypes = [C.c_int] + [C.c_void_p] * 5 + [C.c_ulonglong] * 2 + [C.c_uint]

    def run(which, grid, adj, inp, out, skip, in_step=256, out_step=256):
        p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        g = np.ascontiguousarray(grid, np.uint32)
        gd = np.asarray(g.shape[::-1], np.uint32)
        a = np.ascontiguousarray(adj, np.uint32)
        assert L.emul_stencil4d(which, p(g), p(gd), p(a), p(inp), p(out), in_step, out_step, skip) == 0
    return run


@pytest.mark.parametrize("dims,pa", [((16, 16, 16, 16), False), ((16, 24, 16, 20), True), ((20, 20, 28, 16), False)])
def test_emulated_kernel_bodies_match_oracle_bit_for_bit(built, o4, emul, dims, pa):
    from bricklib_b200 import synth
    dec = built.Decomp4D(dims, 8, pa)
    n = dec.nbricks * 256
    inp = synth.uniform(n, seed=3)
    for s in (0, 1, 2, 3):
        exp = np.full(n, -7.0)
        o4.brick_sweep(dec.adj, o4.sweep_list(dec.grid, s), inp, exp)
        # 0 = line_kernel in its default CTA shape, 1 = gather_kernel (the twin), 41 / 21 / 11 = other shapes
        for which in (0, 1, 41, 21, 11):
            got = np.full(n, -7.0)
            emul(which, dec.grid, dec.adj, inp, got, s)
            assert np.array_equal(got, exp), (s, which)
    dec.close()


def test_emulated_line_kernel_on_a_grid_that_is_not_a_lattice(built, o4, emul):
    """cells permuted: the next cell of a line is not the i+ neighbour, so the marching registers are
    useless and the reload paths run; plus interleaved fields (steps of 512 and 768 elements)"""
    from bricklib_b200 import synth
    dec = built.Decomp4D((16, 24, 16, 20), 8, False)
    grid = np.random.default_rng(5).permutation(dec.grid.reshape(-1)).reshape(dec.grid.shape)
    n = dec.nbricks
    inp = synth.uniform(n * 512, seed=4)
    for s in (0, 2):
        blist = o4.sweep_list(grid, s)
        exp = np.full(n * 768, -7.0)
        o4.brick_sweep(dec.adj, blist, inp[256:], exp[512:], 512, 768)  # field 1 of 2 -> field 2 of 3
        for which in (0, 1):
            got = np.full(n * 768, -7.0)
            emul(which, grid, dec.adj, inp[256:], got[512:], s, 512, 768)
            assert np.array_equal(got, exp), (s, which)
    dec.close()


@pytest.mark.parametrize("flag", ["", "-a", "-P 4", "-P 6 -a"])
def test_driver_validation_pipeline_with_emulated_kernels(built, flag):
    """tests/emul/weak4d_host_check.cpp: everything drivers/weak4d.cpp -v does on the host (decomposition,
    brick fill, the array twin of drivers/host_twin.h, compareBrick<4>) with the device calls replaced by the
    emulated kernel bodies and the exchange by skin -> ghost copies between the emulated ranks of the periodic
    process grid (-P 4: 2x2x1x1, -P 6: 3x2x1x1, where the two neighbours along an axis differ)"""
    out = os.path.join(ROOT, "tests", "emul", "_build")
    os.makedirs(out, exist_ok=True)
    exe = os.path.join(out, "weak4d_host_check")
    src = os.path.join(ROOT, "tests", "emul", "weak4d_host_check.cpp")
    libdir = os.path.join(ROOT, "bricklib_b200", "lib")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fopenmp", "-I" + os.path.join(ROOT, "include"), src, "-o", exe,
                           "-L" + libdir, "-lbricks_b200", "-Wl,-rpath," + libdir])
    r = subprocess.run([exe] + flag.split(), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "match 1" in r.stdout and "match 0" not in r.stdout, r.stdout + r.stderr


# ---------------------------------------------------------------- the kernels on a B200 ----
def _worker(what, *args, timeout=300, variant=None):
    cmd = [sys.executable, os.path.join(ROOT, "tests", "gpu4d_worker.py"), what] + [str(a) for a in args]
    env = dict(os.environ)
    env.pop("BRICKS_4D_VARIANT", None)
    if variant:
        env["BRICKS_4D_VARIANT"] = variant
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=env)
    assert r.returncode == 0, "worker failed (rc %d):\n%s\n%s" % (r.returncode, r.stdout[-2000:], r.stderr[-4000:])
    line = [x for x in r.stdout.splitlines() if x.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [None, "220", "211", "411", "111"])
def test_gpu_4d_kernels_match_oracle(built, variant):
    """None = the library's default (2 x 2 lines per CTA, 256-bit row loads); the others are the CTA shapes
    and the 128-bit load width of profiles/r03_4d/variants.log, selected with BRICKS_4D_VARIANT"""
    res = _worker("parity", variant=variant)
    assert len(res) == 3 * 4 * 2 + 4 + 2
    assert res[-1]["cells"] == 0
    for r in res:
        assert r["relerr"] <= REL_TOL, r
        assert r["untouched_ok"], r


@pytest.mark.gpu
def test_gpu_4d_exchange_period_matches_oracle(built):
    """exchange + ST_ITER = 8 sweeps (weak/main-4d.cpp:247-277), twice, on one periodic rank"""
    res = _worker("period")
    assert res["nregions"] == 312 and res["interior_bricks"] > 0
    assert res["relerr_interior"] <= REL_TOL, res


@pytest.mark.gpu
def test_gpu_4d_full_size_field_matches_oracle(built):
    """a 64^4 subdomain (+ ghost): 160 000 bricks, 41 M points, every one against the oracle"""
    res = _worker("full", 64, timeout=600)
    assert res["relerr"] <= REL_TOL and res["relerr_naive"] <= REL_TOL and res["line_vs_naive"] <= REL_TOL, res
    print("4-D 9-point, one sweep over %d bricks: line %.3f ms (%.0f GB/s), naive %.3f ms" %
          (res["nbricks"], res["line_ms"], res["line_GBps"], res["naive_ms"]))
