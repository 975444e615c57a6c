"""GPU parity of the device-side array <-> brick copies (bricks_copy_array) against the CPU oracle's
restatement of copyToBrick / copyFromBrick (include/bricksetup.h:139-221): byte copies, so the bar
is bit-exact.  Cases: the reference's single-node shape (N + 2*GZ bricks, PADDING = 8), a
BrickDecomp numbering with a skipped ghost band, brick counts along i that are not a multiple of
the eight-brick strip, odd padding (8-byte path) and an interleaved storage (step 1024)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(ctx, built, oracle, dims, pad, ghost, grid, adj, nbricks, step, seed):
    from bricklib_b200 import synth
    d = [int(x) for x in dims]
    ext = [d[a] + 2 * (pad[a] + ghost[a]) for a in range(3)]
    arr = synth.uniform(ext[0] * ext[1] * ext[2], seed=seed).reshape(ext[2], ext[1], ext[0])
    plan = built.Plan(ctx, grid, adj)
    _, grid_dev = plan.device_tables()
    d_arr = ctx.upload(arr)
    d_br = ctx.upload(np.full(nbricks * step, -3.0))
    built.copy_array_dev(ctx, True, d, pad, ghost, d_arr, grid_dev, d_br, step)
    ctx.sync()
    got = ctx.download(d_br, nbricks * step)
    exp = np.full(nbricks * step, -3.0)
    oracle.copy_brick(True, d, pad, ghost, arr, grid, exp, step)
    assert np.array_equal(got, exp), "array -> bricks differs from copyToBrick"
    # and back: only the interior `dims` region of the array is written
    d_back = ctx.upload(np.full(arr.size, -5.0))
    built.copy_array_dev(ctx, False, d, pad, ghost, d_back, grid_dev, d_br, step)
    ctx.sync()
    back = ctx.download(d_back, arr.size).reshape(arr.shape)
    exp_back = np.full(arr.shape, -5.0)
    oracle.copy_brick(False, d, pad, ghost, exp_back, grid, exp, step)
    assert np.array_equal(back, exp_back), "bricks -> array differs from copyFromBrick"
    o = [pad[a] + ghost[a] for a in range(3)]
    inner = back[o[2]:o[2] + d[2], o[1]:o[1] + d[1], o[0]:o[0] + d[0]]
    assert np.array_equal(inner, arr[o[2]:o[2] + d[2], o[1]:o[1] + d[1], o[0]:o[0] + d[0]])
    for p in (d_arr, d_br, d_back):
        ctx.free(p)
    plan.close()


@pytest.mark.parametrize("n,pad", [(32, 8), (64, 8), (40, 3)])
def test_single_node_shape(ctx, built, oracle, n, pad):
    """init_grid numbering over (N + 2*GZ)/8 bricks per axis, whole grid copied (3axis.cpp:20-30)."""
    sg = n + 16
    grid, adj = built.init_grid((sg // 8,) * 3)
    _case(ctx, built, oracle, (sg,) * 3, (pad,) * 3, (0,) * 3, grid, adj, adj.shape[0], 512, 5)


@pytest.mark.parametrize("dims", [(32, 32, 32), (72, 16, 24), (16, 40, 88)])
@pytest.mark.parametrize("step", [512, 1024])
def test_decomp_numbering_skips_the_ghost_band(ctx, built, oracle, dims, step):
    dec = built.Decomp(dims, 8)
    _case(ctx, built, oracle, dims, (8, 8, 8), (8, 8, 8), dec.grid, dec.adj, dec.nbricks, step, 9)


def test_full_size_round_trip(ctx, built):
    """512^3 + ghost: array -> bricks -> array is the identity on the interior, at BASELINE size."""
    from bricklib_b200 import synth
    dec = built.Decomp((512, 512, 512), 8)
    ext = 512 + 16
    arr = synth.uniform(ext ** 3, seed=2).reshape(ext, ext, ext)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    _, grid_dev = plan.device_tables()
    d_arr, d_back = ctx.upload(arr), ctx.malloc(arr.nbytes)
    ctx.memset(d_back, 0, arr.nbytes)
    d_br = ctx.malloc(dec.nbricks * 4096)
    built.copy_array_dev(ctx, True, (512,) * 3, (0,) * 3, (8,) * 3, d_arr, grid_dev, d_br)
    built.copy_array_dev(ctx, False, (512,) * 3, (0,) * 3, (8,) * 3, d_back, grid_dev, d_br)
    ctx.sync()
    back = ctx.download(d_back, arr.size).reshape(arr.shape)
    assert np.array_equal(back[8:-8, 8:-8, 8:-8], arr[8:-8, 8:-8, 8:-8])
    assert not back[:8].any() and not back[:, :8].any() and not back[:, :, -8:].any()
    for p in (d_arr, d_back, d_br):
        ctx.free(p)
    plan.close()


def test_array_layout_baseline_kernel(ctx, built, oracle):
    """bricks_array_stencil7, the device counterpart of d3pt7_arr (stencils/3axis.cu:39-49), against the
    oracle's array form of 7pt.py on a padded array; ragged box sizes, points outside the box untouched."""
    from bricklib_b200 import synth
    from conftest import REL_TOL, relerr
    coeff = synth.coefficients()[:7]
    for ext, lo, hi in [((40, 24, 48), (8, 8, 8), (32, 16, 40)), ((50, 21, 37), (1, 2, 3), (49, 20, 36))]:
        arr = synth.uniform(ext[0] * ext[1] * ext[2], seed=6).reshape(ext[2], ext[1], ext[0])
        d_in, d_out = ctx.upload(arr), ctx.upload(np.full(arr.size, -2.0))
        built.array_stencil7(ctx, d_in, d_out, ext, lo, hi, coeff)
        ctx.sync()
        got = ctx.download(d_out, arr.size).reshape(arr.shape)
        exp = np.full(arr.shape, -2.0)
        oracle.array_sweep(0, arr, exp, (lo[2], hi[2], lo[1], hi[1], lo[0], hi[0]), coeff)
        assert relerr(got, exp) <= REL_TOL
        ctx.free(d_in)
        ctx.free(d_out)
