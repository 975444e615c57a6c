// TEST INFRASTRUCTURE.  The host side of drivers/weak4d.cpp without a GPU: the same decomposition, brick
// fill, array twin (drivers/host_twin.h) and compareBrick<4> the driver uses for -v, with the device calls
// replaced by the host emulation of the kernel bodies (stencil4d_body.cuh) and the exchange of the periodic
// process grid (-P ranks, emulated one after the other) by copies of the skin regions of the rank each
// rank map names onto the ghost regions.  Exit code 0 = on every rank the brick path and the array twin
// agree to 1e-12 after one exchange + ST_ITER sweeps.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "../../bricklib_b200/csrc/stencil4d_body.cuh"
#include "../../drivers/host_twin.h"
#include "bricks/brick-decomp.h"
#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"

typedef Brick<Dim<4, 4, 4, 4>, Dim<4>> Brick4D;
typedef BrickDecomp<4, 4, 4, 4, 4> Decomp4D;

int main(int argc, char **argv) {
  // usage: weak4d_host_check [-a] [-P nranks]
  bool aligned = false;
  int nranks = 1;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "-a")) aligned = true;
    if (!strcmp(argv[i], "-P") && i + 1 < argc) nranks = atoi(argv[++i]);
  }
  const std::vector<unsigned> dom = {16, 24, 16, 20};
  const unsigned GZ = 8, TILE = 4;
  int pdims[4];
  cart_dims_create(nranks, 4, pdims);

  struct RankState {
    std::unique_ptr<Decomp4D> dec;
    BrickInfo<4> *info = nullptr;
    BrickStorage sa, sb;
    bElem *in_ptr = nullptr;
  };
  std::vector<RankState> R(nranks);
  std::vector<long> ext;
  unsigned g[4];
  for (int r = 0; r < nranks; ++r) {
    BricksComm comm = cart_create(nranks, r, 4, nullptr);
    for (int i = 0; i < 4; ++i) comm.dims[i] = pdims[i];
    int coo[4];
    cart_coords(comm, r, coo);
    R[r].dec.reset(new Decomp4D(dom, GZ));
    Decomp4D &dec = *R[r].dec;
    dec.comm = comm;
    dec.page_align = aligned;
    populate(comm, dec, 0, 1, coo);
    std::vector<BitSet> skin;
    allneighbors(0, 1, 4, skin);
    skin.erase(skin.begin() + (skin.size() / 2));
    dec.initialize(skin);
    R[r].info = new BrickInfo<4>(dec.getBrickInfo());
    for (int a = 0; a < 4; ++a) g[a] = dec.gridDim(a);
    ext = {(long)g[0] * TILE, (long)g[1] * TILE, (long)g[2] * TILE, (long)g[3] * TILE};
    R[r].sa = R[r].info->allocate(256);
    R[r].sb = R[r].info->allocate(256);
    memset(R[r].sa.dat.get(), 0, R[r].sa.step * R[r].sa.chunks * 8);
    memset(R[r].sb.dat.get(), 0, R[r].sb.step * R[r].sb.chunks * 8);
    Brick4D bIn(R[r].info, R[r].sa, 0);
    R[r].in_ptr = randomArray(ext, 4321 + r);
    copyToBrick<4>(ext, R[r].in_ptr, const_cast<unsigned *>(dec.gridData()), bIn);
  }

  // BrickDecomp::exchange on the periodic process grid: region i of rank r comes from the skin region i of
  // the rank its rank map names (skin regions are not written by the exchange, so it can run in place)
  for (int r = 0; r < nranks; ++r) {
    Decomp4D &dec = *R[r].dec;
    for (size_t i = 0; i < dec.ghost.size(); ++i) {
      const auto &gr = dec.ghost[i];
      const auto &sk = dec.skin[i];
      const int src = dec.rank_map.at(gr.neighbor.set);
      memcpy(R[r].sa.dat.get() + (size_t)(gr.pos + gr.first_pad) * 256,
             R[src].sa.dat.get() + (size_t)(sk.pos + sk.first_pad) * 256,
             (size_t)(gr.len - gr.first_pad - gr.last_pad) * 256 * 8);
    }
  }
  for (int r = 0; r < nranks; ++r) {
    unsigned *grid = const_cast<unsigned *>(R[r].dec->gridData());
    auto sweep = [&](double *out, const double *in, long skip) {
      s4::Sweep4 p;
      p.grid = grid, p.adj = &R[r].info->adj[0][0], p.in = in, p.out = out, p.in_step = p.out_step = 256;
      p.n0 = g[0], p.n1 = g[1], p.n2 = g[2], p.n3 = g[3], p.s = (unsigned)(skip / TILE);
      const long long nb = (long long)s4::line_blocks(&p, 2, 2);
#pragma omp parallel for
      for (long long b = 0; b < nb; ++b)
        for (unsigned t = 0; t < 256; ++t) s4::line_body<2, 2, false>(p, (unsigned)b, t);
    };
    double *A = R[r].sa.dat.get(), *B = R[r].sb.dat.get();
    sweep(B, A, 0);
    for (int i = 0; i < 3; ++i) {
      sweep(A, B, i * 2 + 1);
      sweep(B, A, i * 2 + 2);
    }
    sweep(A, B, GZ);
  }

  // the array twin, as drivers/weak4d.cpp -v runs it: every ghost band first, then the sweeps
  std::vector<double *> arrays(nranks);
  for (int r = 0; r < nranks; ++r) arrays[r] = R[r].in_ptr;
  const long dml[4] = {(long)dom[0], (long)dom[1], (long)dom[2], (long)dom[3]};
  for (int r = 0; r < nranks; ++r)
    twin::fill_ghost4d(R[r].in_ptr, arrays, dml, GZ, [&](int sl, int sk, int sj, int si) {
      BitSet n;
      if (si) n.flip(si * 1);
      if (sj) n.flip(sj * 2);
      if (sk) n.flip(sk * 3);
      if (sl) n.flip(sl * 4);
      return R[r].dec->rank_map.at(n.set);
    });
  bool all_ok = true;
  for (int r = 0; r < nranks; ++r) {
    bElem *out_ptr = zeroArray(ext);
    bElem *p = R[r].in_ptr, *q = out_ptr;
    for (int it = 0; it < 8; ++it) {
      long lo[4], hi[4];
      for (unsigned a = 0; a < 4; ++a) lo[a] = it + 1, hi[a] = ext[a] - it - 1;
      twin::sweep4d(p, q, ext.data(), lo, hi);
      std::swap(p, q);
    }
    Brick4D bGot(R[r].info, R[r].sa, 0);
    unsigned *grid = const_cast<unsigned *>(R[r].dec->gridData());
    const std::vector<long> dlist(dml, dml + 4), pad0(4, 0), gzv(4, (long)GZ);
    const bool match = compareBrick<4>(dlist, pad0, gzv, p, grid, bGot);
    const double worst = maxRelErrBrick<4>(dlist, pad0, gzv, p, grid, bGot);
    printf("rank %d of %d (%dx%dx%dx%d): match %d worst %.3e regions %zu bricks %u\n", r, nranks, pdims[0], pdims[1],
           pdims[2], pdims[3], (int)match, worst, R[r].dec->ghost.size(), R[r].info->nbricks);
    all_ok = all_ok && match && worst <= 1e-12;
    free(out_ptr);
  }
  return all_ok ? 0 : 1;
}
