"""d3pt27 -- the 27-point stencil with 27 runtime coefficients of stencils/3axis.cpp:165-247
(SURVEY.md 8f item 4).  PINNED: tests/golden/d3pt27_block.npz holds a 16^3 block (and the first
computed 8^3 corner) of the output of the reference's own d3pt27(), run unmodified through
oracle/refshim/ref_3axis.cpp (`python oracle/gen_golden.py d3pt27`), with the input block it depends
on; d3pt27() itself checks its brick form against its array form (compareBrick, 3axis.cpp:236-237).
CPU: the oracle restates the source text (term t at (k + t/9 - 1, j + (t/3)%3 - 1, i + t%3 - 1), term
16 read at (k+1, j+1, i) as BOTH reference forms have it) and reproduces the golden block; its brick
form equals its array form.  GPU: the marching kernel against the golden block and the oracle to
1e-12 relative, and against the naive twin at 512^3."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, REL_TOL, relerr


def golden_block():
    return np.load(os.path.join(GOLDEN, "d3pt27_block.npz"))


@pytest.mark.parametrize("which", ["block", "corner"])
def test_oracle_matches_reference_golden(oracle, which):
    fx = golden_block()
    inp, exp = fx["in_" + which], fx["out_" + which]
    n = exp.shape[0]
    out = np.zeros_like(inp)
    oracle.array_sweep(27, inp, out, (1, n + 1, 1, n + 1, 1, n + 1), fx["coeff"])
    # sums of 27 terms of both signs: error against the scale of the terms (coefficients in
    # [-0.5, 0.5), inputs in [0, 1)), as in test_stencil_cond.py
    scale = float(np.sum(np.abs(fx["coeff"])))
    assert np.max(np.abs(out[1:-1, 1:-1, 1:-1] - exp)) <= 1e-14 * scale
    # the quirk is real: reading term 16 at (k, j+1, i) instead does NOT reproduce the reference
    plain = np.zeros_like(exp)
    for t in range(27):
        dk, dj, di = t // 9 - 1, (t // 3) % 3 - 1, t % 3 - 1
        plain += fx["coeff"][t] * inp[1 + dk:1 + dk + n, 1 + dj:1 + dj + n, 1 + di:1 + di + n]
    assert np.max(np.abs(plain - exp)) > 1e-3


def test_oracle_point_set_follows_the_reference_text(oracle):
    from bricklib_b200 import synth
    coeff = synth.coefficients()[:27]
    off, c = oracle.points(27, coeff)
    assert len(off) == 27 and np.array_equal(c, coeff)
    for t in range(27):
        exp = (t % 3 - 1, (t // 3) % 3 - 1, 1 if t == 16 else t // 9 - 1)
        assert tuple(int(x) for x in off[t]) == exp, t
    # (k, j+1, i) is never read; (k+1, j+1, i) is read twice (terms 16 and 25)
    pts = [tuple(int(x) for x in o) for o in off]
    assert (0, 1, 0) not in pts and pts.count((0, 1, 1)) == 2


def test_oracle_brick_form_equals_array_form(built, oracle):
    from bricklib_b200 import synth
    n, pad = 24, 8
    sg = n + 16
    grid, adj = built.init_grid((sg // 8,) * 3)
    ext = sg + 2 * pad
    arr = synth.uniform(ext ** 3, seed=4).reshape(ext, ext, ext)
    coeff = synth.coefficients()[:27]
    bdat = np.zeros(adj.shape[0] * 512)
    oracle.copy_brick(True, (sg,) * 3, (pad,) * 3, (0,) * 3, arr, grid, bdat)
    interior = grid[1:-1, 1:-1, 1:-1].reshape(-1)
    bout = np.zeros_like(bdat)
    oracle.brick_sweep(27, adj, interior, bdat, bout, 512, 512, coeff)
    aout = np.zeros_like(arr)
    lo, hi = pad + 8, pad + 8 + n
    oracle.array_sweep(27, arr, aout, (lo, hi, lo, hi, lo, hi), coeff)
    back = np.zeros_like(arr)
    oracle.copy_brick(False, (sg,) * 3, (pad,) * 3, (0,) * 3, back, grid, bout)
    assert relerr(back[lo:hi, lo:hi, lo:hi], aout[lo:hi, lo:hi, lo:hi]) <= 1e-14


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(16, 16, 16), (32, 32, 32), (16, 24, 40), (48, 16, 24)])
def test_gpu_27pt_matches_oracle(ctx, built, oracle, dims):
    from bricklib_b200 import synth
    dec = built.Decomp(dims, 8)
    inp = synth.uniform(dec.nbricks * 512, seed=1)
    coeff = synth.coefficients()[:27]
    plan = built.Plan(ctx, dec.grid, dec.adj)
    d_in, d_out = ctx.upload(inp), ctx.upload(np.full(dec.nbricks * 512, -7.0))
    plan.stencil(built.ST_27PT_COEFF, d_in, d_out, coeff=coeff)
    ctx.sync()
    got = ctx.download(d_out, dec.nbricks * 512)
    exp = np.full(dec.nbricks * 512, -7.0)
    oracle.brick_sweep(27, dec.adj, dec.grid.reshape(-1), inp, exp, 512, 512, coeff)
    assert relerr(got, exp) <= REL_TOL
    assert np.all(got[:512] == -7.0)  # the sentinel brick is outside the plan
    # interleaved storage (single-node drivers): in at offset 0, out at offset 512 of step 1024
    st = np.zeros(dec.nbricks * 1024)
    st.reshape(-1, 1024)[:, :512] = inp.reshape(-1, 512)
    d_st = ctx.upload(st)
    plan.stencil(built.ST_27PT_COEFF, d_st, d_st + 512 * 8, 1024, 1024, coeff=coeff)
    ctx.sync()
    got2 = ctx.download(d_st, dec.nbricks * 1024).reshape(-1, 1024)[:, 512:]
    assert relerr(got2[1:], exp.reshape(-1, 512)[1:]) <= REL_TOL
    for p in (d_in, d_out, d_st):
        ctx.free(p)
    plan.close()


@pytest.mark.gpu
def test_gpu_27pt_matches_reference_golden_block(ctx, built, oracle):
    """the reference's 18^3 input block placed in a 16^3 subdomain with its one-brick ghost ring (a
    32^3 array, the block's halo in the ghost bricks): the owned 16^3 points equal the reference's
    own output block"""
    fx = golden_block()
    dec = built.Decomp((16, 16, 16), 8)
    t = dec.grid.shape
    arr = np.zeros((t[0] * 8, t[1] * 8, t[2] * 8))
    arr[7:25, 7:25, 7:25] = fx["in_block"]
    field = np.zeros(dec.nbricks * 512)
    oracle.copy_brick(True, [t[2] * 8, t[1] * 8, t[0] * 8], [0, 0, 0], [0, 0, 0], arr, dec.grid, field)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    d_in, d_out = ctx.upload(field), ctx.upload(np.zeros(dec.nbricks * 512))
    plan.stencil(built.ST_27PT_COEFF, d_in, d_out, coeff=fx["coeff"])
    ctx.sync()
    got = ctx.download(d_out, dec.nbricks * 512)
    back = np.zeros_like(arr)
    oracle.copy_brick(False, [t[2] * 8, t[1] * 8, t[0] * 8], [0, 0, 0], [0, 0, 0], back, dec.grid, got)
    scale = float(np.sum(np.abs(fx["coeff"])))
    assert np.max(np.abs(back[8:24, 8:24, 8:24] - fx["out_block"])) <= REL_TOL * scale
    for p in (d_in, d_out):
        ctx.free(p)
    plan.close()


@pytest.mark.gpu
def test_gpu_27pt_full_size_against_the_naive_twin(ctx, built):
    from bricklib_b200 import synth
    dec = built.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    coeff = synth.coefficients()[:27]
    d_a = ctx.upload(built.uniform(nb * 512, seed=3))
    d_b, d_c = ctx.malloc(nb * 4096), ctx.malloc(nb * 4096)
    ctx.memset(d_b, 0, nb * 4096)
    ctx.memset(d_c, 0, nb * 4096)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    plan.stencil(built.ST_27PT_COEFF, d_a, d_b, coeff=coeff)
    plan.stencil_naive(built.ST_27PT_COEFF, d_a, d_c, coeff=coeff)
    ctx.sync()
    assert relerr(ctx.download(d_b, nb * 512), ctx.download(d_c, nb * 512)) <= REL_TOL
    for p in (d_a, d_b, d_c):
        ctx.free(p)
    plan.close()
