// TEST INFRASTRUCTURE.  The host side of drivers/weak4d.cpp without a GPU: the same decomposition, brick
// fill, array twin (drivers/host_twin.h) and compareBrick<4> the driver uses for -v, with the device calls
// replaced by the host emulation of the kernel bodies (stencil4d_body.cuh) and the single-rank periodic
// exchange replaced by a copy of every skin region onto its ghost region.  Exit code 0 = the brick path and
// the array twin agree to 1e-12 after one exchange + ST_ITER sweeps.
#include <cstdio>
#include <cstring>

#include "../../bricklib_b200/csrc/stencil4d_body.cuh"
#include "../../drivers/host_twin.h"
#include "bricks/brick-decomp.h"
#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"

typedef Brick<Dim<4, 4, 4, 4>, Dim<4>> Brick4D;
typedef BrickDecomp<4, 4, 4, 4, 4> Decomp4D;

int main(int argc, char **argv) {
  std::vector<unsigned> dom = {16, 24, 16, 20};
  const bool aligned = argc > 1 && !strcmp(argv[1], "-a");
  const unsigned GZ = 8, TILE = 4;
  BricksComm comm = cart_create(1, 0, 4, nullptr);
  int coo[4] = {0, 0, 0, 0};
  Decomp4D dec(dom, GZ);
  dec.comm = comm;
  dec.page_align = aligned;
  populate(comm, dec, 0, 1, coo);
  std::vector<BitSet> skin;
  allneighbors(0, 1, 4, skin);
  skin.erase(skin.begin() + (skin.size() / 2));
  dec.initialize(skin);
  BrickInfo<4> info = dec.getBrickInfo();
  const unsigned g[4] = {dec.gridDim(0), dec.gridDim(1), dec.gridDim(2), dec.gridDim(3)};
  unsigned *grid = const_cast<unsigned *>(dec.gridData());
  BrickStorage sa = info.allocate(256), sb = info.allocate(256);
  memset(sa.dat.get(), 0, sa.step * sa.chunks * 8);
  memset(sb.dat.get(), 0, sb.step * sb.chunks * 8);
  Brick4D bIn(&info, sa, 0);
  const std::vector<long> ext = {(long)g[0] * TILE, (long)g[1] * TILE, (long)g[2] * TILE, (long)g[3] * TILE};
  bElem *in_ptr = randomArray(ext, 4321);
  copyToBrick<4>(ext, in_ptr, grid, bIn);

  // exchange on one periodic rank: every neighbour is this rank
  for (size_t i = 0; i < dec.ghost.size(); ++i) {
    const auto &gr = dec.ghost[i];
    const auto &sk = dec.skin[i];
    memcpy(sa.dat.get() + (size_t)(gr.pos + gr.first_pad) * 256, sa.dat.get() + (size_t)(sk.pos + sk.first_pad) * 256,
           (size_t)(gr.len - gr.first_pad - gr.last_pad) * 256 * 8);
  }
  auto sweep = [&](double *out, const double *in, long skip) {
    s4::Sweep4 p;
    p.grid = grid, p.adj = &info.adj[0][0], p.in = in, p.out = out, p.in_step = p.out_step = 256;
    p.n0 = g[0], p.n1 = g[1], p.n2 = g[2], p.n3 = g[3], p.s = (unsigned)(skip / TILE);
    const long long nb = (long long)s4::line_blocks(&p, 2, 2);
#pragma omp parallel for
    for (long long b = 0; b < nb; ++b)
      for (unsigned t = 0; t < 256; ++t) s4::line_body<2, 2, false>(p, (unsigned)b, t);
  };
  double *A = sa.dat.get(), *B = sb.dat.get();
  sweep(B, A, 0);
  for (int i = 0; i < 3; ++i) {
    sweep(A, B, i * 2 + 1);
    sweep(B, A, i * 2 + 2);
  }
  sweep(A, B, GZ);

  // the array twin, as drivers/weak4d.cpp -v runs it
  std::vector<double *> arrays = {in_ptr};
  const long dml[4] = {(long)dom[0], (long)dom[1], (long)dom[2], (long)dom[3]};
  twin::fill_ghost4d(in_ptr, arrays, dml, GZ, [&](int sl, int sk, int sj, int si) {
    BitSet n;
    if (si) n.flip(si * 1);
    if (sj) n.flip(sj * 2);
    if (sk) n.flip(sk * 3);
    if (sl) n.flip(sl * 4);
    return dec.rank_map.at(n.set);
  });
  bElem *out_ptr = zeroArray(ext);
  bElem *p = in_ptr, *q = out_ptr;
  for (int it = 0; it < 8; ++it) {
    long lo[4], hi[4];
    for (unsigned a = 0; a < 4; ++a) lo[a] = it + 1, hi[a] = ext[a] - it - 1;
    twin::sweep4d(p, q, ext.data(), lo, hi);
    std::swap(p, q);
  }
  const std::vector<long> dlist(dml, dml + 4), pad0(4, 0), gzv(4, (long)GZ);
  const bool match = compareBrick<4>(dlist, pad0, gzv, p, grid, bIn);
  const double worst = maxRelErrBrick<4>(dlist, pad0, gzv, p, grid, bIn);
  printf("match %d worst %.3e regions %zu bricks %u\n", (int)match, worst, dec.ghost.size(), info.nbricks);
  return match && worst <= 1e-12 ? 0 : 1;
}
