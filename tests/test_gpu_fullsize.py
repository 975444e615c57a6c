"""Full-size (512^3, the size BASELINE.json names) checks of the CUDA path: the WHOLE field (all
287 496 bricks, ghost bricks included) against the compiled reference (oracle/_ref: the reference's
own codegen brick kernels, a fraction of a second per sweep on the host cores), a sample of bricks
against the slow C oracle as a second opinion, and size-independent properties."""
import numpy as np
import pytest

from conftest import REL_TOL, STENCILS, relerr

pytestmark = pytest.mark.gpu

COEFF_SUM = {7: 0.4 + 6 * 0.1, 13: 0.4 + 6 * (0.07 + 0.03), 25: 0.1 + 6 * (0.06 + 0.045 + 0.03 + 0.015)}


def _c125_sum():
    C = [0.1, 0.04, 0.03, 0.01, 0.006, 0.004, 0.005, 0.002, 0.003, 0.001]
    cls = {(0, 0, 0): 0, (0, 0, 1): 1, (0, 0, 2): 2, (0, 1, 1): 3, (0, 1, 2): 4, (0, 2, 2): 5, (1, 1, 1): 6,
           (1, 1, 2): 7, (1, 2, 2): 8, (2, 2, 2): 9}
    s = 0.0
    for a in range(-2, 3):
        for b in range(-2, 3):
            for c in range(-2, 3):
                s += C[cls[tuple(sorted((abs(a), abs(b), abs(c))))]]
    return s


@pytest.fixture(scope="module")
def big(ctx, built):
    from bricklib_b200 import synth
    dec = built.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    plan = built.Plan(ctx, dec.grid, dec.adj)
    d_a = ctx.malloc(nb * 4096)
    d_b = ctx.malloc(nb * 4096)
    d_c = ctx.malloc(nb * 4096)
    fa = built.uniform(nb * 512, seed=31)
    fb = built.uniform(nb * 512, seed=32)
    yield dict(dec=dec, plan=plan, d=(d_a, d_b, d_c), fa=fa, fb=fb, nb=nb)
    plan.close()
    for p in (d_a, d_b, d_c):
        ctx.free(p)
    dec.close()


@pytest.mark.parametrize("which", STENCILS)
def test_512_whole_field_matches_compiled_reference(ctx, built, big, which):
    """every brick of the 66^3-brick grid, CUDA sweep vs the reference's generated brick kernel
    (brick(...) of weak/main.cpp:140-170 / stencils/3axis.cpp:57-66 compiled from /root/reference by
    oracle/Makefile) on the same field: <= 1e-12 max relative error over all 147 197 952 points"""
    from oracle import refload as R
    from bricklib_b200 import synth
    if not R.available():
        pytest.skip("oracle/_ref is not built for this host ISA")
    dec, plan, (d_a, d_b, _), nb = big["dec"], big["plan"], big["d"], big["nb"]
    coeff = synth.coefficients() if which == 0 else None
    ctx.h2d(d_a, big["fa"])
    ctx.memset(d_b, 0, nb * 4096)
    plan.stencil(which, d_a, d_b, 512, 512, coeff)
    ctx.sync()
    got = ctx.download(d_b, nb * 512)
    exp = R.aligned_empty(nb * 512)
    exp[:] = 0.0
    inp = R.aligned_empty(nb * 512)
    inp[:] = big["fa"]
    blist = dec.grid.reshape(-1)
    R.brick_sweep_rm(which, dec.adj, blist, inp, exp, 512, coeff)
    assert blist.size == 66 ** 3 and np.unique(blist).size == blist.size
    assert relerr(got[512:], exp[512:]) <= REL_TOL
    assert np.all(got[:512] == 0.0)  # the sentinel brick is never written


@pytest.mark.parametrize("which", STENCILS)
def test_512_sampled_bricks_match_oracle(ctx, built, oracle, big, which):
    from bricklib_b200 import synth
    dec, plan, (d_a, d_b, _), nb = big["dec"], big["plan"], big["d"], big["nb"]
    coeff = synth.coefficients() if which == 0 else None
    ctx.h2d(d_a, big["fa"])
    ctx.memset(d_b, 0, nb * 4096)
    plan.stencil(which, d_a, d_b, 512, 512, coeff)
    ctx.sync()
    got = ctx.download(d_b, nb * 512).reshape(nb, 512)
    g = dec.grid
    rng = np.random.default_rng(which)
    sample = [g[0, 0, 0], g[65, 65, 65], g[0, 65, 33], g[1, 1, 1], g[64, 64, 64], g[1, 32, 64], g[33, 0, 1]]
    sample += list(rng.choice(g.reshape(-1), 600 if which != 125 else 300, replace=False))
    sample = np.unique(np.array(sample, np.uint32))
    exp = np.zeros(nb * 512)
    oracle.brick_sweep(which, dec.adj, sample, big["fa"], exp, 512, 512, coeff)
    assert relerr(got[sample], exp.reshape(nb, 512)[sample]) <= REL_TOL
    assert np.all(got[0] == 0.0)  # the sentinel brick is never written


@pytest.mark.parametrize("which", (7, 13, 25, 125))
def test_512_constant_field_gives_coefficient_sum(ctx, built, big, which):
    """stencil(c) = c * sum(coefficients) wherever no sentinel/garbage brick is read, i.e. on the
    interior bricks (the region the reference compares, weak/main.cu:321-323)."""
    dec, plan, (d_a, d_b, _), nb = big["dec"], big["plan"], big["d"], big["nb"]
    ctx.h2d(d_a, np.full(nb * 512, 1.5))
    plan.stencil(which, d_a, d_b, 512, 512)
    ctx.sync()
    got = ctx.download(d_b, nb * 512).reshape(nb, 512)
    interior = dec.grid[1:-1, 1:-1, 1:-1].reshape(-1)
    s = COEFF_SUM.get(which) or _c125_sum()
    assert relerr(got[interior], np.full((interior.size, 512), 1.5 * s)) <= REL_TOL


@pytest.mark.parametrize("which", (7, 125))
def test_512_linearity(ctx, built, big, which):
    """S(a + 2b) = S(a) + 2 S(b) over the whole 66^3-brick field, to rounding."""
    plan, (d_a, d_b, d_c), nb = big["plan"], big["d"], big["nb"]
    fa, fb = big["fa"], big["fb"]
    outs = []
    for f in (fa, fb, fa + 2.0 * fb):
        ctx.h2d(d_a, f)
        plan.stencil(which, d_a, d_b, 512, 512)
        ctx.sync()
        outs.append(ctx.download(d_b, nb * 512))
    lhs, rhs = outs[2][512:], (outs[0] + 2.0 * outs[1])[512:]
    assert relerr(lhs, rhs) <= 1e-13 * 50


def test_512_sweep_is_deterministic_and_checksum_stable(ctx, built, big):
    """Two launches over the same input give bit-identical fields (no atomics, no races)."""
    plan, (d_a, d_b, d_c), nb = big["plan"], big["d"], big["nb"]
    ctx.h2d(d_a, big["fa"])
    plan.stencil(125, d_a, d_b, 512, 512)
    plan.stencil(125, d_a, d_c, 512, 512)
    ctx.sync()
    x = ctx.download(d_b, nb * 512)
    y = ctx.download(d_c, nb * 512)
    assert np.array_equal(x[512:], y[512:])
