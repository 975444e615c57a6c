"""Pins the oracle: every table and value it produces is checked against tests/golden/*,
which oracle/gen_golden.py dumped from the UNMODIFIED compiled reference (oracle/_ref), and
against oracle/_ref itself when that library is present on this machine."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, REL_TOL, STENCILS, fnv1a64, relerr


def _check_tables(dec, rec):
    assert dec.nbricks == rec["nbricks"]
    assert [int(x) for x in dec.sep_pos] == rec["sep_pos"]
    assert [int(x) for x in dec.skin_size] == rec["skin_size"]
    for name, tab in (("ghost", dec.ghost), ("skin", dec.skin)):
        assert len(tab) == len(rec[name]["pos"])
        for k, v in rec[name].items():
            assert np.array_equal(tab[k].astype(np.uint64), np.array(v, np.uint64)), (name, k)


@pytest.mark.parametrize("idx", range(5))
def test_oracle_decomp_tables_match_reference(oracle, golden_decomp, idx):
    rec = golden_decomp["decomp"][idx]
    dec = oracle.Decomp(rec["dims"], 8)
    _check_tables(dec, rec)
    if rec["nbricks"] <= 6000:  # hashing 7.7 M adjacency entries in Python is slow; 512^3 goes via _ref below
        assert fnv1a64(dec.grid) == rec["grid_fnv"]
        assert fnv1a64(dec.adj) == rec["adj_fnv"]


def test_oracle_rank_map_matches_reference(oracle, golden_decomp):
    for case in golden_decomp["decomp"][0]["rank_map"]:
        rm = oracle.rank_map(case["procdims"], case["coo"])
        assert rm == {int(k, 16): v for k, v in case["map"].items()}


@pytest.mark.parametrize("dims", [(16, 16, 16), (16, 24, 32)])
def test_oracle_stencils_match_reference_outputs(oracle, dims):
    """brick sweep (codegen brick kernel of the reference) and array sweep (ST_CPU) fixtures."""
    from bricklib_b200 import synth
    fx = np.load(os.path.join(GOLDEN, "stencil_%dx%dx%d.npz" % dims))
    dec = oracle.Decomp(dims, 8)
    inp = synth.uniform(dec.nbricks * 512, seed=11)
    coeff = synth.coefficients()
    interior = fx["interior_bricks"]
    assert np.array_equal(interior, dec.grid[1:-1, 1:-1, 1:-1].reshape(-1))
    for w in STENCILS:
        cf = coeff if w == 0 else None
        out = np.zeros(dec.nbricks * 512)
        oracle.brick_sweep(w, dec.adj, dec.grid.reshape(-1), inp, out, 512, 512, cf)
        got = out.reshape(dec.nbricks, 512)[interior]
        assert relerr(got, fx["brick_%d" % w]) <= 1e-14, w
        # array twin: the same field laid out as a padded array, interior box only
        t = dec.grid.shape
        arr = np.zeros((t[0] * 8, t[1] * 8, t[2] * 8))
        oracle.copy_brick(False, [t[2] * 8, t[1] * 8, t[0] * 8], [0, 0, 0], [0, 0, 0], arr, dec.grid, inp)
        aout = np.zeros_like(arr)
        oracle.array_sweep(w, arr, aout, (8, arr.shape[0] - 8, 8, arr.shape[1] - 8, 8, arr.shape[2] - 8), cf)
        assert relerr(aout[8:-8, 8:-8, 8:-8], fx["array_%d" % w]) <= 1e-14, w


def test_oracle_brick_and_array_forms_agree(oracle):
    """The reference's own self-check (compareBrick of array result vs brick result,
    stencils/3axis.cpp:78-86), at the strict tolerance."""
    from bricklib_b200 import synth
    dec = oracle.Decomp((24, 16, 32), 8)
    inp = synth.uniform(dec.nbricks * 512, seed=5)
    t = dec.grid.shape
    ext = [t[2] * 8, t[1] * 8, t[0] * 8]
    arr = np.zeros((ext[2], ext[1], ext[0]))
    oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], arr, dec.grid, inp)
    for w in STENCILS:
        cf = synth.coefficients() if w == 0 else None
        bout = np.zeros_like(inp)
        oracle.brick_sweep(w, dec.adj, dec.grid[1:-1, 1:-1, 1:-1].reshape(-1), inp, bout, 512, 512, cf)
        aout = np.zeros_like(arr)
        oracle.array_sweep(w, arr, aout, (8, ext[2] - 8, 8, ext[1] - 8, 8, ext[0] - 8), cf)
        back = np.zeros_like(arr)
        oracle.copy_brick(False, ext, [0, 0, 0], [0, 0, 0], back, dec.grid, bout)
        assert relerr(back[8:-8, 8:-8, 8:-8], aout[8:-8, 8:-8, 8:-8]) <= REL_TOL


def test_oracle_against_compiled_reference_when_present(oracle):
    """Direct check against oracle/_ref (present wherever build() ran next to /root/reference,
    and on the GPU box as a prebuilt file).  Skips where it cannot load."""
    from oracle import refload as R
    if not R.available():
        pytest.skip("oracle/_ref not built for this host ISA")
    from bricklib_b200 import synth
    ref = R.RefDecomp((512, 512, 512), 8)
    dec = oracle.Decomp((512, 512, 512), 8)
    assert np.array_equal(ref.grid, dec.grid)
    assert np.array_equal(ref.adj, dec.adj)
    assert np.array_equal(ref.sep_pos, dec.sep_pos)
    ref.close()
    small = oracle.Decomp((32, 24, 16), 8)
    inp = synth.uniform(small.nbricks * 512, seed=2)
    for w in STENCILS:
        cf = synth.coefficients() if w == 0 else None
        a, b = np.zeros_like(inp), np.zeros_like(inp)
        oracle.brick_sweep(w, small.adj, small.grid.reshape(-1), inp, a, 512, 512, cf)
        R.brick_sweep_rm(w, small.adj, small.grid.reshape(-1), inp, b, 512, cf)
        assert relerr(a, b) <= 1e-14


def test_oracle_exchange_is_a_periodic_shift(oracle):
    """After one emulated exchange on a 2x1x1 periodic grid, every ghost brick of rank r holds
    the skin brick of its neighbour that sits at the same global position."""
    dec = oracle.Decomp((16, 16, 16), 8)
    P = [2, 1, 1]
    st = []
    for r in range(2):
        s = np.zeros(dec.nbricks * 512)
        # tag each brick with (rank, brick index)
        s.reshape(dec.nbricks, 512)[:] = (r * 100000 + np.arange(dec.nbricks))[:, None]
        st.append(s)
    oracle.exchange(st, dec, P)
    g = dec.grid
    for r in range(2):
        f = st[r].reshape(dec.nbricks, 512)[:, 0]
        # ghost cell (k,j,i) of rank r = interior cell wrapped by the subdomain extent (2 bricks)
        # of the rank that owns it on the periodic grid
        for k in range(4):
            for j in range(4):
                for i in range(4):
                    inside = all(1 <= c <= 2 for c in (k, j, i))
                    if inside:
                        assert f[g[k, j, i]] == r * 100000 + g[k, j, i]
                        continue
                    wk, wj, wi = ((c - 1) % 2 + 1 for c in (k, j, i))
                    # process grid dims are (2,1,1) with index 0 slowest = k axis
                    owner = (r + (0 if 1 <= k <= 2 else 1)) % 2
                    assert f[g[k, j, i]] == owner * 100000 + g[wk, wj, wi], (r, k, j, i)
