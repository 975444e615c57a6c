/*
 * oracle/refshim/ref_decomp.cpp -- TEST INFRASTRUCTURE, not product code.
 *
 * Thin extern "C" window onto the UNMODIFIED reference decomposition
 * (include/brick-mpi.h:178-460 BrickDecomp::initialize, :730-753 populate,
 * src/brick-mpi.cpp:9-52 allneighbors / skin3d_*), compiled from the sources
 * where they lie under /root/reference by oracle/Makefile into oracle/_ref/.
 * It exists so that tests and oracle/gen_golden.py can pin the oracle and the
 * product's BrickDecomp against the real thing.  Nothing here is shipped.
 */
#include <cstdint>
#include <cstring>
#include <vector>

#include <brick-mpi.h>

int refstub_cart_ndims = 3;
int refstub_cart_dims[8] = {1, 1, 1, 1, 1, 1, 1, 1};

namespace {
typedef BrickDecomp<3, 8, 8, 8> Decomp3;

struct Handle {
  Decomp3 *dec;
  unsigned tdims[3];
  unsigned nbricks;
  unsigned (*adj)[27];
};

const std::vector<BitSet> &pick_skin(int which) {
  if (which == 1) return skin3d_normal;
  if (which == 2) return skin3d_bad;
  return skin3d_good;
}
}  // namespace

extern "C" {

void *ref_decomp_new(const unsigned *dims, unsigned gz, int skin) {
  Handle *h = new Handle;
  h->dec = new Decomp3({dims[0], dims[1], dims[2]}, gz);
  h->dec->comm = 0;
  h->dec->initialize(pick_skin(skin));
  BrickInfo<3> bi = h->dec->getBrickInfo();
  h->nbricks = bi.nbricks;
  h->adj = bi.adj;
  for (int d = 0; d < 3; ++d) h->tdims[d] = dims[d] / 8 + 2 * (gz / 8);
  return h;
}

void ref_decomp_free(void *p) {
  Handle *h = (Handle *)p;
  free(h->adj);
  delete h->dec;
  delete h;
}

unsigned ref_decomp_nbricks(void *p) { return ((Handle *)p)->nbricks; }

void ref_decomp_sep_pos(void *p, unsigned *out) {
  Handle *h = (Handle *)p;
  for (int i = 0; i < 3; ++i) out[i] = h->dec->sep_pos[i];
}

unsigned ref_decomp_num_regions(void *p) { return (unsigned)((Handle *)p)->dec->ghost.size(); }

/* which: 0 = ghost[], 1 = skin[]; seven parallel output arrays of num_regions entries */
void ref_decomp_regions(void *p, int which, uint64_t *neighbor, unsigned *skin_st, unsigned *skin_ed,
                        unsigned *pos, unsigned *len, unsigned *first_pad, unsigned *last_pad) {
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

unsigned ref_decomp_num_skin(void *p) { return (unsigned)((Handle *)p)->dec->skin_size.size(); }

void ref_decomp_skin_size(void *p, long *out) {
  Handle *h = (Handle *)p;
  for (size_t i = 0; i < h->dec->skin_size.size(); ++i) out[i] = h->dec->skin_size[i];
}

/* grid in [k][j][i] order over the ghost-including brick grid (operator[] of the reference) */
void ref_decomp_grid(void *p, unsigned *out) {
  Handle *h = (Handle *)p;
  size_t n = 0;
  for (unsigned k = 0; k < h->tdims[2]; ++k)
    for (unsigned j = 0; j < h->tdims[1]; ++j)
      for (unsigned i = 0; i < h->tdims[0]; ++i) out[n++] = (*h->dec)[k][j][i];
}

void ref_decomp_adj(void *p, unsigned *out) {
  Handle *h = (Handle *)p;
  std::memcpy(out, h->adj, sizeof(unsigned) * 27 * (size_t)h->nbricks);
}

/* rank_map after populate(cart, decomp, 0, 1, coo) on a periodic procdims grid;
 * keys/ranks come back in allneighbors() order (27 entries). */
void ref_decomp_rank_map(void *p, const int *procdims, const int *coo_in, uint64_t *keys, int *ranks) {
  Handle *h = (Handle *)p;
  refstub_set_cart(3, procdims);
  int coo[3] = {coo_in[0], coo_in[1], coo_in[2]};
  MPI_Comm comm = 0;
  h->dec->rank_map.clear();
  populate(comm, *h->dec, 0, 1, coo);
  std::vector<BitSet> nb;
  allneighbors(0, 1, 3, nb);
  for (size_t i = 0; i < nb.size(); ++i) {
    keys[i] = nb[i].set;
    ranks[i] = h->dec->rank_map[nb[i].set];
  }
}

}  // extern "C"
