"""GPU parity tests proper: the CUDA stencil kernels, called through the C ABI
(bricks_plan_create / bricks_stencil), against the CPU oracle on identical seeded inputs,
against the golden outputs of the compiled reference, and -- at the full 512^3 size of
BASELINE.json -- through size-independent properties plus an oracle check of sampled bricks.
Bar: max relative error <= 1e-12 in fp64 (north_star)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, REL_TOL, STENCILS, relerr

pytestmark = pytest.mark.gpu


def _run(ctx, capi, plan, which, inp, nb, coeff=None, in_step=512, out_step=512, in_off=0, out_off=0, naive=False):
    d_in = ctx.upload(inp)
    out_elems = nb * out_step
    d_out = ctx.upload(np.full(out_elems, -7.0))
    fn = plan.stencil_naive if naive else plan.stencil
    fn(which, d_in + in_off * 8, d_out + out_off * 8, in_step, out_step, coeff)
    ctx.sync()
    got = ctx.download(d_out, out_elems)
    ctx.free(d_in)
    ctx.free(d_out)
    return got


@pytest.mark.parametrize("dims", [(16, 16, 16), (32, 32, 32), (16, 24, 40), (48, 16, 24)])
@pytest.mark.parametrize("which", STENCILS)
def test_stencil_matches_oracle_on_decomposed_grid(ctx, built, oracle, dims, which):
    from bricklib_b200 import synth
    dec = built.Decomp(dims, 8)
    inp = synth.uniform(dec.nbricks * 512, seed=1)
    coeff = synth.coefficients() if which == 0 else None
    plan = built.Plan(ctx, dec.grid, dec.adj)
    got = _run(ctx, built, plan, which, inp, dec.nbricks, coeff)
    exp = np.full(dec.nbricks * 512, -7.0)
    oracle.brick_sweep(which, dec.adj, dec.grid.reshape(-1), inp, exp, 512, 512, coeff)
    assert relerr(got, exp) <= REL_TOL
    # bricks outside the plan (only the sentinel here) are untouched
    assert np.all(got[:512] == -7.0)
    plan.close()


@pytest.mark.parametrize("which", STENCILS)
def test_stencil_matches_reference_golden_outputs(ctx, built, which):
    """Directly against the outputs of the reference's generated brick kernels."""
    from bricklib_b200 import synth
    for dims in [(16, 16, 16), (16, 24, 32)]:
        fx = np.load(os.path.join(GOLDEN, "stencil_%dx%dx%d.npz" % dims))
        dec = built.Decomp(dims, 8)
        inp = synth.uniform(dec.nbricks * 512, seed=11)
        coeff = synth.coefficients() if which == 0 else None
        plan = built.Plan(ctx, dec.grid, dec.adj)
        got = _run(ctx, built, plan, which, inp, dec.nbricks, coeff)
        interior = fx["interior_bricks"]
        assert relerr(got.reshape(-1, 512)[interior], fx["brick_%d" % which]) <= REL_TOL
        plan.close()


@pytest.mark.parametrize("which", STENCILS)
def test_box_plans_odd_widths_and_untouched_bricks(ctx, built, oracle, which):
    """Plans over several boxes (odd i-widths leave a half-empty brick pair; thin k-slabs give
    one-layer columns): bricks in the boxes match the oracle, all others keep their fill."""
    from bricklib_b200 import synth
    dec = built.Decomp((32, 32, 32), 8)
    boxes = [((1, 1, 1), (4, 5, 5)), ((0, 0, 0), (6, 6, 1)), ((5, 0, 1), (6, 6, 6)), ((0, 5, 5), (5, 6, 6))]
    blist = np.concatenate([dec.grid[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]].reshape(-1) for lo, hi in boxes])
    inp = synth.uniform(dec.nbricks * 512, seed=4)
    coeff = synth.coefficients() if which == 0 else None
    plan = built.Plan(ctx, dec.grid, dec.adj, boxes)
    got = _run(ctx, built, plan, which, inp, dec.nbricks, coeff)
    exp = np.full(dec.nbricks * 512, -7.0)
    oracle.brick_sweep(which, dec.adj, blist, inp, exp, 512, 512, coeff)
    assert relerr(got, exp) <= REL_TOL
    untouched = np.setdiff1d(np.arange(dec.nbricks), blist)
    assert np.all(got.reshape(-1, 512)[untouched] == -7.0)
    plan.close()


@pytest.mark.parametrize("which", (0, 125))
def test_single_node_layout_interleaved_fields(ctx, built, oracle, which):
    """The single-node shape of the reference (stencils/3axis.cu:77-198): init_grid numbering
    (out-of-range neighbours alias brick 0), ONE storage with bIn at offset 0 and bOut at offset
    512 (step 1024), interior bricks only."""
    from bricklib_b200 import synth
    sb = 6  # STRIDEB for N = 32: (32 + 2*8) / 8
    grid, adj = built.init_grid((sb, sb, sb))
    nb = adj.shape[0]
    st = synth.uniform(nb * 1024, seed=8)
    coeff = synth.coefficients() if which == 0 else None
    plan = built.Plan(ctx, grid, adj, [((1, 1, 1), (sb - 1, sb - 1, sb - 1))])
    d = ctx.upload(st)
    plan.stencil(which, d, d + 512 * 8, 1024, 1024, coeff)
    ctx.sync()
    got = ctx.download(d, nb * 1024)
    ctx.free(d)
    exp = st.copy()
    blist = grid[1:-1, 1:-1, 1:-1].reshape(-1)
    oracle.brick_sweep(which, adj, blist, exp, exp[512:], 1024, 1024, coeff)
    assert relerr(got, exp) <= REL_TOL
    assert np.array_equal(got.reshape(nb, 1024)[:, :512], st.reshape(nb, 1024)[:, :512])  # input intact
    plan.close()


def test_batched_fields_share_one_plan(ctx, built, oracle):
    """strong-scaling shape (strong/main.cu:85-99,464-469): many (in,out) storage pairs, one
    BrickInfo, one launch."""
    from bricklib_b200 import synth
    dec = built.Decomp((16, 16, 16), 8)
    nsub = 5
    plan = built.Plan(ctx, dec.grid, dec.adj)
    ins = [synth.uniform(dec.nbricks * 512, seed=20 + s) for s in range(nsub)]
    d_in = [ctx.upload(x) for x in ins]
    d_out = [ctx.upload(np.zeros(dec.nbricks * 512)) for _ in range(nsub)]
    n0 = built.launch_count()
    plan.stencil(built.ST_125PT, d_in, d_out, 512, 512)
    ctx.sync()
    assert built.launch_count() - n0 == 1
    for s in range(nsub):
        got = ctx.download(d_out[s], dec.nbricks * 512)
        exp = np.zeros(dec.nbricks * 512)
        oracle.brick_sweep(125, dec.adj, dec.grid.reshape(-1), ins[s], exp, 512, 512)
        assert relerr(got, exp) <= REL_TOL
    for p in d_in + d_out:
        ctx.free(p)
    plan.close()


@pytest.mark.parametrize("which", STENCILS)
def test_naive_twin_agrees(ctx, built, oracle, which):
    from bricklib_b200 import synth
    dec = built.Decomp((16, 16, 24), 8)
    inp = synth.uniform(dec.nbricks * 512, seed=6)
    coeff = synth.coefficients() if which == 0 else None
    plan = built.Plan(ctx, dec.grid, dec.adj)
    a = _run(ctx, built, plan, which, inp, dec.nbricks, coeff, naive=True)
    b = _run(ctx, built, plan, which, inp, dec.nbricks, coeff)
    assert relerr(b[512:], a[512:]) <= REL_TOL
    plan.close()


def test_bad_arguments_return_errors(ctx, built):
    dec = built.Decomp((16, 16, 16), 8)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    d = ctx.malloc(dec.nbricks * 4096)
    with pytest.raises(built.BricksError):
        plan.stencil(99, d, d + 4096, 512, 512)  # unknown stencil (and aliasing-free pointers)
    with pytest.raises(built.BricksError):
        plan.stencil(7, d, d, 512, 512)  # in == out
    with pytest.raises(built.BricksError):
        plan.stencil(0, d, d + 4096, 512, 512, None)  # 7pt.py needs coefficients
    with pytest.raises(built.BricksError):
        built.Plan(ctx, dec.grid, dec.adj, [((0, 0, 0), (99, 1, 1))])  # box outside the grid
    ctx.free(d)
    plan.close()
