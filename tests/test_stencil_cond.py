"""cond -- stencils/cond.py:23-30, the max/abs stencil the reference's single drivers run as d3cond /
d3condcu (stencils/3axis.cpp:94-163, single/cuda.cpp:23): out = | sum_t coeff[t] * max(in_t, 0) | over
7pt.py's point set.  PINNED: tests/golden/cond_*.npz holds the outputs of the reference's own codegen
brick kernel and of its array form (oracle/gen_golden.py cond, from oracle/_ref).

Inputs and coefficients carry both signs (synth.uniform - 0.5), otherwise max() and |.| would be
no-ops.  With mixed signs the sum cancels, so the error is measured against the scale of the terms,
not against a result that may be arbitrarily close to zero: |got - exp| <= 1e-12 * max(|exp|, S) with
S = sum_t |coeff[t]| * max|in| (about 1); for results of ordinary size this is the 1e-12 relative bar
of the other stencils."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

COND = 1
DIMS = [(16, 16, 16), (16, 24, 32)]


def cond_inputs(nelem):
    from bricklib_b200 import synth
    return synth.uniform(nelem, seed=13) - 0.5, synth.coefficients()[:7] - 0.5


def scaled_err(got, exp, coeff):
    scale = float(np.sum(np.abs(coeff))) * 0.5
    return float(np.max(np.abs(got - exp) / np.maximum(np.abs(exp), scale)))


@pytest.mark.parametrize("dims", DIMS)
def test_oracle_matches_reference_golden(oracle, dims):
    fx = np.load(os.path.join(GOLDEN, "cond_%dx%dx%d.npz" % dims))
    dec = oracle.Decomp(dims, 8)
    inp, coeff = cond_inputs(dec.nbricks * 512)
    interior = fx["interior_bricks"]
    assert np.array_equal(interior, dec.grid[1:-1, 1:-1, 1:-1].reshape(-1))
    out = np.zeros(dec.nbricks * 512)
    oracle.brick_sweep(COND, dec.adj, dec.grid.reshape(-1), inp, out, 512, 512, coeff)
    got = out.reshape(dec.nbricks, 512)[interior]
    assert scaled_err(got, fx["brick_1"], coeff) <= 1e-14
    assert np.all(got >= 0) and np.any(inp < 0)
    t = dec.grid.shape
    arr = np.zeros((t[0] * 8, t[1] * 8, t[2] * 8))
    oracle.copy_brick(False, [t[2] * 8, t[1] * 8, t[0] * 8], [0, 0, 0], [0, 0, 0], arr, dec.grid, inp)
    aout = np.zeros_like(arr)
    oracle.array_sweep(COND, arr, aout, (8, arr.shape[0] - 8, 8, arr.shape[1] - 8, 8, arr.shape[2] - 8), coeff)
    assert scaled_err(aout[8:-8, 8:-8, 8:-8], fx["array_1"], coeff) <= 1e-14


def test_max_and_abs_both_matter(oracle):
    """the fixture inputs exercise both non-linearities: dropping either changes the result"""
    dec = oracle.Decomp((16, 16, 16), 8)
    inp, coeff = cond_inputs(dec.nbricks * 512)
    blist = dec.grid[1:-1, 1:-1, 1:-1].reshape(-1)
    cond, lin, relu_only = (np.zeros(dec.nbricks * 512) for _ in range(3))
    oracle.brick_sweep(COND, dec.adj, blist, inp, cond, 512, 512, coeff)
    oracle.brick_sweep(0, dec.adj, blist, inp, lin, 512, 512, coeff)
    oracle.brick_sweep(0, dec.adj, blist, np.maximum(inp, 0.0), relu_only, 512, 512, coeff)
    assert not np.allclose(cond, np.abs(lin))          # max() matters
    assert np.any(relu_only < 0)                        # |.| matters
    assert np.allclose(cond, np.abs(relu_only), rtol=0, atol=1e-15)


# ------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("dims", DIMS + [(32, 32, 32), (48, 16, 24)])
def test_gpu_cond_matches_oracle_and_golden(ctx, built, oracle, dims):
    dec = built.Decomp(dims, 8)
    inp, coeff = cond_inputs(dec.nbricks * 512)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    d_in, d_out = ctx.upload(inp), ctx.upload(np.full(dec.nbricks * 512, -7.0))
    plan.stencil(built.ST_COND, d_in, d_out, coeff=coeff)
    ctx.sync()
    got = ctx.download(d_out, dec.nbricks * 512)
    exp = np.full(dec.nbricks * 512, -7.0)
    oracle.brick_sweep(COND, dec.adj, dec.grid.reshape(-1), inp, exp, 512, 512, coeff)
    assert scaled_err(got, exp, coeff) <= 1e-12
    assert np.all(got[:512] == -7.0)
    path = os.path.join(GOLDEN, "cond_%dx%dx%d.npz" % dims)
    if os.path.exists(path):  # directly against the reference's codegen brick kernel
        fx = np.load(path)
        assert scaled_err(got.reshape(-1, 512)[fx["interior_bricks"]], fx["brick_1"], coeff) <= 1e-12
    for p in (d_in, d_out):
        ctx.free(p)
    plan.close()


@pytest.mark.gpu
def test_gpu_cond_full_size_against_the_naive_twin(ctx, built):
    dec = built.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    inp, coeff = cond_inputs(nb * 512)
    d_a = ctx.upload(inp)
    d_b, d_c = ctx.malloc(nb * 4096), ctx.malloc(nb * 4096)
    ctx.memset(d_b, 0, nb * 4096)
    ctx.memset(d_c, 0, nb * 4096)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    plan.stencil(built.ST_COND, d_a, d_b, coeff=coeff)
    plan.stencil_naive(built.ST_COND, d_a, d_c, coeff=coeff)
    ctx.sync()
    got, exp = ctx.download(d_b, nb * 512), ctx.download(d_c, nb * 512)
    assert scaled_err(got, exp, coeff) <= 1e-12 and np.all(got >= 0)
    for p in (d_a, d_b, d_c):
        ctx.free(p)
    plan.close()
