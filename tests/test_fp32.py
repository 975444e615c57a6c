"""fp32 bricks: the reference's element type is overridable (`#define bElem float`,
include/vecscatter.h:12-14).  Pinned against the reference's own generated brick kernels built with
-DbElem=float through its code generator (oracle/_ref, symbols reff<N>_*; AVX-512 hosts).

Parity bar restated for fp32: 1e-5 max relative error.  The fp64 bar (1e-12) is ~5e3 ulp of headroom over
the measured 2e-15; the same terms summed in float in a different order differ by up to ~n * 2^-24 for n
positive terms of similar size (125 * 6e-8 = 7.5e-6), measured 1e-6 between the reference's vector code and a
scalar left-to-right sum."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import ROOT

REL_TOL_F32 = 1e-5
STENCILS_F32 = (7, 13, 25, 125)


def relerr32(got, exp):
    got, exp = np.asarray(got, np.float64), np.asarray(exp, np.float64)
    return float(np.max(np.abs(got - exp) / np.maximum(np.abs(exp), 1e-30)))


def _f32_field(built, n, seed):
    return built.uniform(n, seed=seed).astype(np.float32)


@pytest.mark.parametrize("which", STENCILS_F32)
def test_f32_oracle_matches_the_compiled_float_reference(oracle, which):
    from oracle import refload as R
    if not R.available_f32():
        pytest.skip("oracle/_ref has no fp32 build for this host (AVX-512 only)")
    from bricklib_b200 import synth
    dec = oracle.Decomp((32, 24, 16), 8)
    n = dec.nbricks * 512
    x = synth.uniform(n, seed=12).astype(np.float32)
    blist = dec.grid.reshape(-1)
    a, b = np.zeros(n, np.float32), np.zeros(n, np.float32)
    oracle.brick_sweep_f32(which, dec.adj, blist, x, a)
    R.brick_sweep_rm_f32(which, dec.adj, blist, x, b, 512)
    assert a.any() and relerr32(a, b) <= REL_TOL_F32
    # and fp32 is fp64 rounded, not something else
    d = np.zeros(n)
    oracle.brick_sweep(which, dec.adj, blist, x.astype(np.float64), d)
    assert relerr32(b, d) <= REL_TOL_F32


def test_host_headers_compile_for_float_bricks():
    """`#define bElem float` in front of the drop-in headers: Brick / BrickStorage / BrickDecomp / copyToBrick /
    BrickStencilPlan::run (-> bricks_stencil_f32) all follow the element type"""
    src = r'''
#define bElem float
#include "bricks/brick-b200.h"
#include "bricks/brick-decomp.h"
#include "bricks/bricksetup.h"
#include "bricks/brickcompare.h"
#include "bricks/multiarray.h"
typedef Brick<Dim<8, 8, 8>, Dim<1>> Brick3D;
int main() {
  static_assert(sizeof(bElem) == 4, "fp32 build");
  BrickDecomp<3, 8, 8, 8> dec({32, 32, 32}, 8);
  dec.initialize(skin3d_good);
  BrickInfo<3> info = dec.getBrickInfo();
  BrickStorage a = info.allocate(Brick3D::BRICKSIZE), b = info.allocate(Brick3D::BRICKSIZE);
  Brick3D bin(&info, a, 0), bout(&info, b, 0);
  const unsigned g[3] = {dec.gridDim(0), dec.gridDim(1), dec.gridDim(2)};
  BrickStencilPlan::Box box = {{0, 0, 0}, {g[0], g[1], g[2]}};
  BrickStencilPlan plan(nullptr, dec.gridData(), g, &info.adj[0][0], info.nbricks, {box});
  plan.run(BRICKS_ST_7PT, bin, bout);
  ExchangeView ev = dec.exchangeView(a);
  return (int)sizeof(bin.dat[0]);
}
'''
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "f32_tu.cpp")
        open(p, "w").write(src)
        r = subprocess.run(["/usr/bin/g++", "-std=c++17", "-fsyntax-only", "-fopenmp", "-I" + os.path.join(ROOT, "include"), p],
                           capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("which", STENCILS_F32)
@pytest.mark.parametrize("dims", [(16, 16, 16), (32, 24, 40)])
def test_f32_sweep_matches_oracle_and_naive_twin(ctx, built, oracle, which, dims):
    dec = built.Decomp(dims, 8)
    nb = dec.nbricks
    x = _f32_field(built, nb * 512, 21)
    d_in = ctx.upload(x)
    d_out = ctx.upload(np.zeros(nb * 512, np.float32))
    d_nv = ctx.upload(np.zeros(nb * 512, np.float32))
    plan = built.Plan(ctx, dec.grid, dec.adj)
    plan.stencil_f32(which, d_in, d_out)
    plan.stencil_naive_f32(which, d_in, d_nv)
    ctx.sync()
    got = ctx.download(d_out, nb * 512, np.float32)
    twin = ctx.download(d_nv, nb * 512, np.float32)
    exp = np.zeros(nb * 512, np.float32)
    oracle.brick_sweep_f32(which, dec.adj, dec.grid.reshape(-1), x, exp)
    assert exp.any() and np.all(got[:512] == 0.0)  # the sentinel brick is never written
    assert relerr32(got[512:], exp[512:]) <= REL_TOL_F32
    assert relerr32(twin[512:], exp[512:]) <= REL_TOL_F32
    # a second launch on other storages of the same plan (cached tensor maps are per storage pair)
    plan.stencil_f32(which, d_out, d_nv)
    ctx.sync()
    exp2 = np.zeros(nb * 512, np.float32)
    oracle.brick_sweep_f32(which, dec.adj, dec.grid.reshape(-1), got, exp2)
    assert relerr32(ctx.download(d_nv, nb * 512, np.float32)[512:], exp2[512:]) <= REL_TOL_F32
    plan.close()
    for p in (d_in, d_out, d_nv):
        ctx.free(p)
    dec.close()


@pytest.mark.gpu
@pytest.mark.parametrize("which", STENCILS_F32)
def test_f32_512_whole_field_matches_the_compiled_float_reference(ctx, built, which):
    """all 287 496 bricks of the 512^3 subdomain, CUDA float kernel vs the reference's generated float kernel"""
    from oracle import refload as R
    if not R.available_f32():
        pytest.skip("oracle/_ref has no fp32 build for this host (AVX-512 only)")
    dec = built.Decomp((512, 512, 512), 8)
    nb = dec.nbricks
    x = _f32_field(built, nb * 512, 31)
    d_in = ctx.upload(x)
    d_out = ctx.upload(np.zeros(nb * 512, np.float32))
    plan = built.Plan(ctx, dec.grid, dec.adj)
    plan.stencil_f32(which, d_in, d_out)
    ctx.sync()
    got = ctx.download(d_out, nb * 512, np.float32)
    exp = np.zeros(nb * 512, np.float32)
    R.brick_sweep_rm_f32(which, dec.adj, dec.grid.reshape(-1), x, exp, 512)
    assert relerr32(got[512:], exp[512:]) <= REL_TOL_F32
    plan.close()
    ctx.free(d_in)
    ctx.free(d_out)
    dec.close()


@pytest.mark.gpu
def test_f32_storage_through_the_exchange_and_two_sweeps(ctx, built, oracle):
    """the exchange moves bytes: a float storage is exchanged with offsets scaled by its element size
    (step = 256 eight-byte units for 512 floats), bit-exact against the oracle's emulation on the same
    bytes; two float sweeps on top agree with the float oracle"""
    dims, which = (32, 40, 24), 13
    dec = built.Decomp(dims, 8)
    dec.rank_map([1, 1, 1], [0, 0, 0])
    nb = dec.nbricks
    ex = built.Exchange(ctx, 0, *dec.exchange_plan(256))
    x = _f32_field(built, nb * 512, 41)
    d_a = ctx.upload(x)
    d_b = ctx.upload(np.zeros(nb * 512, np.float32))
    ex.run(d_a, built.STREAM_COMM)
    ctx.sync()
    got = ctx.download(d_a, nb * 512, np.float32)
    odec = oracle.Decomp(dims, 8)
    exp = oracle.exchange([x.view(np.float64).copy()], odec, [1, 1, 1], step=256)[0].view(np.float32)
    assert np.array_equal(got.view(np.uint32), exp.view(np.uint32)) and not np.array_equal(got, x)
    plan = built.Plan(ctx, dec.grid, dec.adj)
    plan.stencil_f32(which, d_a, d_b)
    plan.stencil_f32(which, d_b, d_a)
    ctx.sync()
    res = ctx.download(d_a, nb * 512, np.float32)
    t1, t2 = np.zeros(nb * 512, np.float32), np.zeros(nb * 512, np.float32)
    oracle.brick_sweep_f32(which, odec.adj, odec.grid.reshape(-1), exp, t1)
    oracle.brick_sweep_f32(which, odec.adj, odec.grid.reshape(-1), t1, t2)
    interior = odec.grid[1:-1, 1:-1, 1:-1].reshape(-1)
    assert relerr32(res.reshape(nb, 512)[interior], t2.reshape(nb, 512)[interior]) <= 2 * REL_TOL_F32
    plan.close()
    ex.close()
    ctx.free(d_a)
    ctx.free(d_b)
    dec.close()
