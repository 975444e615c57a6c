/*
 * oracle/refshim/ref_4d.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * extern "C" window onto the UNMODIFIED reference 4-D path of weak/main-4d.cpp:
 *   - BrickDecomp<4, 4,4,4,4>(dom, GZ = 8).initialize(allneighbors minus the centre)
 *     (weak/main-4d.cpp:98-106, include/brick-mpi.h:322-460), built twice by oracle/Makefile:
 *     page-aligned (what weak/CMakeLists.txt builds for cpu-4d; 2 KiB bricks -> regions padded
 *     to whole pages) and with -DDECOMP_PAGEUNALIGN (what the CUDA drivers use);
 *   - the codegen brick kernel `brick("mpi9pt.py", VSVEC, (4,4,4,4), (VFOLD), b)` of
 *     weak/main-4d.cpp:62-75 with the vector settings that file selects on an AVX2 host
 *     (VSVEC "AVX2", fold 2,2; its `__AVX512__` test never fires, :36-58).
 * The array form of weak/main-4d.cpp:171-193 is a lambda inside main() and cannot be linked; the oracle
 * restates it (oracle/stencil_oracle.c: oracle_array_sweep_4d).
 * Compiled into oracle/_ref/libref4d.so (ref4d_* stencils, ref4a_* page-aligned numbering) and
 * oracle/_ref/libref4d_unalign.so (ref4u_* dense numbering): two libraries because the two
 * instantiations of BrickDecomp<4,...>::initialize would collapse into one symbol in a single link.
 */
#include <omp.h>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include <brick-mpi.h>
#include <brick.h>

#define R4_CAT2(a, b) a##b
#define R4_CAT(a, b) R4_CAT2(a, b)

#ifdef REF4_DECOMP
/* ------------------------------------------------------------------ numbering ---- */
#ifdef DECOMP_PAGEUNALIGN
#define R4SYM(name) R4_CAT(ref4u_, name)
#else
#define R4SYM(name) R4_CAT(ref4a_, name)
#endif
/* the two globals the MPI stub declares */
int refstub_cart_ndims = 4;
int refstub_cart_dims[8] = {1, 1, 1, 1, 1, 1, 1, 1};

namespace {
typedef BrickDecomp<4, 4, 4, 4, 4> Decomp4;
struct Handle {
  Decomp4 *dec;
  unsigned tdims[4];
  unsigned nbricks;
  unsigned (*adj)[81];
};
}  // namespace

extern "C" {

void *R4SYM(new)(const unsigned *dims, unsigned gz) {
  Handle *h = new Handle;
  h->dec = new Decomp4({dims[0], dims[1], dims[2], dims[3]}, gz);
  h->dec->comm = 0;
  std::vector<BitSet> skin;
  allneighbors(0, 1, 4, skin);
  skin.erase(skin.begin() + (skin.size() / 2));
  h->dec->initialize(skin);
  BrickInfo<4> bi = h->dec->getBrickInfo();
  h->nbricks = bi.nbricks;
  h->adj = bi.adj;
  for (int d = 0; d < 4; ++d) h->tdims[d] = dims[d] / 4 + 2 * (gz / 4);
  return h;
}

void R4SYM(free)(void *p) {
  Handle *h = (Handle *)p;
  free(h->adj);
  delete h->dec;
  delete h;
}

unsigned R4SYM(nbricks)(void *p) { return ((Handle *)p)->nbricks; }
void R4SYM(sep_pos)(void *p, unsigned *out) {
  for (int i = 0; i < 3; ++i) out[i] = ((Handle *)p)->dec->sep_pos[i];
}
unsigned R4SYM(num_regions)(void *p) { return (unsigned)((Handle *)p)->dec->ghost.size(); }
unsigned R4SYM(num_skin)(void *p) { return (unsigned)((Handle *)p)->dec->skin_size.size(); }

void R4SYM(regions)(void *p, int which, uint64_t *neighbor, unsigned *skin_st, unsigned *skin_ed, unsigned *pos,
                    unsigned *len, unsigned *first_pad, unsigned *last_pad) {
  Handle *h = (Handle *)p;
  auto &v = which ? h->dec->skin : h->dec->ghost;
  for (size_t i = 0; i < v.size(); ++i) {
    neighbor[i] = v[i].neighbor.set;
    skin_st[i] = v[i].skin_st;
    skin_ed[i] = v[i].skin_ed;
    pos[i] = v[i].pos;
    len[i] = v[i].len;
    first_pad[i] = v[i].first_pad;
    last_pad[i] = v[i].last_pad;
  }
}

void R4SYM(skin_size)(void *p, long *out) {
  Handle *h = (Handle *)p;
  for (size_t i = 0; i < h->dec->skin_size.size(); ++i) out[i] = h->dec->skin_size[i];
}

/* grid in [l][k][j][i] order over the ghost-including brick grid */
void R4SYM(grid)(void *p, unsigned *out) {
  Handle *h = (Handle *)p;
  size_t n = 0;
  for (unsigned l = 0; l < h->tdims[3]; ++l)
    for (unsigned k = 0; k < h->tdims[2]; ++k)
      for (unsigned j = 0; j < h->tdims[1]; ++j)
        for (unsigned i = 0; i < h->tdims[0]; ++i) out[n++] = (*h->dec)[l][k][j][i];
}

/* adjacency rows of the bricks named in the grid only: the reference leaves the rows of the reserved
 * and padding slots uninitialised (malloc), so those are returned as zeros */
void R4SYM(adj)(void *p, unsigned *out) {
  Handle *h = (Handle *)p;
  std::memset(out, 0, sizeof(unsigned) * 81 * (size_t)h->nbricks);
  size_t cells = (size_t)h->tdims[0] * h->tdims[1] * h->tdims[2] * h->tdims[3];
  std::vector<unsigned> g(cells);
  R4SYM(grid)(p, g.data());
  for (size_t c = 0; c < cells; ++c) std::memcpy(out + (size_t)g[c] * 81, h->adj[g[c]], sizeof(unsigned) * 81);
}

/* rank_map after populate(cart, decomp, 0, 1, coo) on a periodic 4-D process grid; 81 entries in
 * allneighbors() order */
void R4SYM(rank_map)(void *p, const int *procdims, const int *coo_in, uint64_t *keys, int *ranks) {
  Handle *h = (Handle *)p;
  refstub_set_cart(4, procdims);
  int coo[4] = {coo_in[0], coo_in[1], coo_in[2], coo_in[3]};
  MPI_Comm comm = 0;
  h->dec->rank_map.clear();
  populate(comm, *h->dec, 0, 1, coo);
  std::vector<BitSet> nb;
  allneighbors(0, 1, 4, nb);
  for (size_t i = 0; i < nb.size(); ++i) {
    keys[i] = nb[i].set;
    ranks[i] = h->dec->rank_map[nb[i].set];
  }
}

}  // extern "C"

#else
/* ------------------------------------------------------------------- stencils ---- */
#include <immintrin.h>
#define R4_VSVEC "AVX2"
#define R4_BDIM 4, 4, 4, 4
#define R4_VFOLD 2, 2
typedef Brick<Dim<R4_BDIM>, Dim<R4_VFOLD>> Brick4D;

namespace {

BrickStorage wrap(bElem *p, long chunks, long step) {
  BrickStorage s;
  s.dat = std::shared_ptr<bElem>(p, [](bElem *) {});
  s.chunks = chunks;
  s.step = (size_t)step;
  return s;
}

struct InfoView {
  BrickInfo<4> info;
  void *scratch;
  InfoView(const unsigned *adj, unsigned nbricks) : info(0) {
    scratch = info.adj;
    info.adj = (BrickInfo<4>::adjlist)adj;
    info.nbricks = nbricks;
  }
  ~InfoView() { free(scratch); }
};

/* offset of element (l,k,j,i) inside one brick in this build's fold order, from the reference accessor */
void fold_table(unsigned *tab) {
  unsigned adj1[81];
  for (int i = 0; i < 81; ++i) adj1[i] = 0;
  InfoView iv(adj1, 1);
  static bElem buf[256];
  BrickStorage st = wrap(buf, 1, 256);
  Brick4D b(&iv.info, st, 0);
  for (unsigned l = 0; l < 4; ++l)
    for (unsigned k = 0; k < 4; ++k)
      for (unsigned j = 0; j < 4; ++j)
        for (unsigned i = 0; i < 4; ++i) tab[((l * 4 + k) * 4 + j) * 4 + i] = (unsigned)(&b[0][l][k][j][i] - buf);
}

void sweep_folded(const unsigned *adj, unsigned nbricks, const unsigned *blist, long nlist, bElem *in_dat,
                  bElem *out_dat, long step) {
  InfoView iv(adj, nbricks);
  BrickStorage sin = wrap(in_dat, nbricks, step), sout = wrap(out_dat, nbricks, step);
  Brick4D in(&iv.info, sin, 0), out(&iv.info, sout, 0);
#pragma omp parallel for schedule(static)
  for (long n = 0; n < nlist; ++n) {
    unsigned b = blist[n];
    brick(REF_PY, R4_VSVEC, (R4_BDIM), (R4_VFOLD), b);
  }
}

}  // namespace

extern "C" {

void ref4d_fold_table(unsigned *tab256) { fold_table(tab256); }
int ref4d_max_threads() { return omp_get_max_threads(); }

/* One sweep of the codegen kernel over blist[0..nlist); storages are row-major [l][k][j][i] bricks
 * (the device layout), converted to and from the build's fold order through fold_table.  Bricks that
 * are not in blist keep their previous contents in out_rm. */
void ref4d_brick_sweep_rm(const unsigned *adj, unsigned nbricks, const unsigned *blist, long nlist,
                          const bElem *in_rm, bElem *out_rm, long step) {
  unsigned tab[256];
  fold_table(tab);
  size_t tot = (size_t)nbricks * step;
  bElem *fin = (bElem *)aligned_alloc(64, tot * sizeof(bElem));
  bElem *fout = (bElem *)aligned_alloc(64, tot * sizeof(bElem));
  long nfield = step / 256;
#pragma omp parallel for
  for (long c = 0; c < (long)nbricks * nfield; ++c)
    for (unsigned e = 0; e < 256; ++e) {
      fin[c * 256 + tab[e]] = in_rm[c * 256 + e];
      fout[c * 256 + tab[e]] = out_rm[c * 256 + e];
    }
  sweep_folded(adj, nbricks, blist, nlist, fin, fout, step);
#pragma omp parallel for
  for (long c = 0; c < (long)nbricks * nfield; ++c)
    for (unsigned e = 0; e < 256; ++e) out_rm[c * 256 + e] = fout[c * 256 + tab[e]];
  free(fin);
  free(fout);
}

/* `reps` ping-pong sweeps in the build's own fold order (what the reference times), after one warm-up;
 * seconds per sweep.  a and b are only read and written in folded order, so any initial contents do. */
double ref4d_brick_sweep_timed(const unsigned *adj, unsigned nbricks, const unsigned *blist, long nlist, bElem *a,
                               bElem *b, long step, int reps) {
  sweep_folded(adj, nbricks, blist, nlist, a, b, step);
  double t0 = omp_get_wtime();
  for (int r = 0; r < reps; ++r) {
    if (r & 1)
      sweep_folded(adj, nbricks, blist, nlist, b, a, step);
    else
      sweep_folded(adj, nbricks, blist, nlist, a, b, step);
  }
  return (omp_get_wtime() - t0) / reps;
}

}  // extern "C"
#endif
