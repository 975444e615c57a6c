// Flat C wrappers over the C++ host templates (include/bricks_b200_host.h).  Host only.
#include "../../include/bricks_b200_host.h"

#include <cstring>
#include <exception>
#include <memory>

#include "../../include/bricks/brick-decomp.h"
#include "../../include/bricks/brickcompare.h"
#include "../../include/bricks/bricksetup.h"
#include "../../include/bricks/multiarray.h"
#include "../../include/bricks/strong-plan.h"
#include "../../include/bricks/zmort.h"

namespace bricks {
void set_error(const char *fmt, ...);
}

typedef BrickDecomp<3, 8, 8, 8> Decomp3;
typedef Brick<Dim<8, 8, 8>, Dim<8>> Brick3;
typedef BrickDecomp<4, 4, 4, 4, 4> Decomp4;
typedef Brick<Dim<4, 4, 4, 4>, Dim<4>> Brick4; /* fold <4> = row-major [l][k][j][i] */
static_assert(Brick4::row_major(), "4-D bricks of the C ABI are row-major");

struct bricks_decomp {
  Decomp3 *dec;
};
struct bricks_decomp4d {
  Decomp4 *dec;
};
struct bricks_strong_plan {
  StrongPlan *plan;
};

namespace {

template <typename F>
int guarded(F f) {
  try {
    f();
    return 0;
  } catch (const std::exception &e) {
    bricks::set_error("%s", e.what());
    return -5;
  }
}

template <typename Region>
void flatten(const std::vector<Region> &v, bricks_region *out) {
  for (size_t i = 0; i < v.size(); ++i) {
    out[i].neighbor = v[i].neighbor.set;
    out[i].skin_st = v[i].skin_st;
    out[i].skin_ed = v[i].skin_ed;
    out[i].pos = v[i].pos;
    out[i].len = v[i].len;
    out[i].first_pad = v[i].first_pad;
    out[i].last_pad = v[i].last_pad;
  }
}

/* a non-owning Brick view over caller memory */
struct View {
  BrickInfo<3> info;
  void *own_adj;
  BrickStorage st;
  View(const unsigned *adj, unsigned nbricks, double *dat, size_t step) : info(1) {
    own_adj = info.adj;
    info.adj = (BrickInfo<3>::adjlist) const_cast<unsigned *>(adj);
    info.nbricks = nbricks;
    st.dat = std::shared_ptr<bElem>(dat, [](bElem *) {});
    st.chunks = nbricks;
    st.step = step;
  }
  ~View() { free(own_adj); }
};

}  // namespace

extern "C" {

int bricks_decomp_create(const unsigned dims[3], unsigned depth, int skin_kind, bricks_decomp **out) {
  if (!dims || !out || depth % 8 || dims[0] % 8 || dims[1] % 8 || dims[2] % 8 || skin_kind < 0 || skin_kind > 2) {
    bricks::set_error("bricks_decomp_create: bad argument (extents and depth must be multiples of 8)");
    return -1;
  }
  for (int a = 0; a < 3; ++a)
    if (dims[a] < 2 * depth) {
      bricks::set_error("bricks_decomp_create: extent %u is smaller than two ghost depths", dims[a]);
      return -1;
    }
  return guarded([&] {
    std::unique_ptr<Decomp3> d(new Decomp3({dims[0], dims[1], dims[2]}, depth));
    d->initialize(skin_kind == 1 ? skin3d_normal : (skin_kind == 2 ? skin3d_bad : skin3d_good));
    *out = new bricks_decomp{d.release()};
  });
}

int bricks_decomp_destroy(bricks_decomp *d) {
  if (d) {
    delete d->dec;
    delete d;
  }
  return 0;
}

int bricks_decomp_sizes(const bricks_decomp *d, unsigned *nbricks, unsigned sep_pos[3], unsigned gdims[3],
                        int *nregions, int *nskin) {
  if (!d) return -1;
  if (nbricks) *nbricks = d->dec->sep_pos[2];
  for (int a = 0; a < 3; ++a) {
    if (sep_pos) sep_pos[a] = d->dec->sep_pos[a];
    if (gdims) gdims[a] = d->dec->gridDim(a);
  }
  if (nregions) *nregions = (int)d->dec->ghost.size();
  if (nskin) *nskin = (int)d->dec->skinlist.size();
  return 0;
}

int bricks_decomp_tables(const bricks_decomp *d, unsigned *grid, unsigned *adj, bricks_region *ghost,
                         bricks_region *skin, long *skin_size) {
  if (!d) return -1;
  Decomp3 &D = *d->dec;
  const size_t cells = (size_t)D.gridDim(0) * D.gridDim(1) * D.gridDim(2);
  if (grid) memcpy(grid, D.gridData(), cells * sizeof(unsigned));
  if (adj) memcpy(adj, D.getBrickInfo().adj, (size_t)D.sep_pos[2] * 27 * sizeof(unsigned));
  if (ghost) flatten(D.ghost, ghost);
  if (skin) flatten(D.skin, skin);
  if (skin_size)
    for (size_t i = 0; i < D.skin_size.size(); ++i) skin_size[i] = D.skin_size[i];
  return 0;
}

int bricks_decomp_rank_map(bricks_decomp *d, const int procdims[3], const int coo_in[3], uint64_t keys[27],
                           int ranks[27]) {
  if (!d || !procdims || !coo_in) return -1;
  BricksComm c;
  c.ndims = 3;
  c.size = 1;
  for (int a = 0; a < 3; ++a) {
    c.dims[a] = procdims[a];
    c.size *= procdims[a];
  }
  int coo[3] = {coo_in[0], coo_in[1], coo_in[2]};
  c.rank = cart_rank(c, coo);
  c.ctx = d->dec->comm.ctx;
  d->dec->comm = c;
  d->dec->rank_map.clear();
  populate(c, *d->dec, 0, 1, coo);
  if (keys && ranks) {
    std::vector<BitSet> nb;
    allneighbors(0, 1, 3, nb);
    for (size_t i = 0; i < nb.size(); ++i) {
      keys[i] = nb[i].set;
      ranks[i] = d->dec->rank_map[nb[i].set];
    }
  }
  return 0;
}

int bricks_decomp_exchange_plan(bricks_decomp *d, size_t step, size_t *recv_off, size_t *send_off, size_t *nbytes,
                                int *recv_peer, int *send_peer) {
  if (!d || !recv_off || !send_off || !nbytes || !recv_peer || !send_peer) return -1;
  Decomp3 &D = *d->dec;
  if (D.rank_map.empty()) {
    bricks::set_error("bricks_decomp_exchange_plan: call bricks_decomp_rank_map first");
    return -1;
  }
  const size_t sb = step * sizeof(double);
  for (size_t i = 0; i < D.ghost.size(); ++i) {
    const auto &g = D.ghost[i];
    const auto &s = D.skin[i];
    recv_off[i] = (size_t)(g.pos + g.first_pad) * sb;
    send_off[i] = (size_t)(s.pos + s.first_pad) * sb;
    nbytes[i] = (size_t)(g.len - g.first_pad - g.last_pad) * sb;
    recv_peer[i] = D.rank_map[g.neighbor.set];
    send_peer[i] = D.rank_map[s.neighbor.set];
  }
  return 0;
}

int bricks_decomp_shift_plan(bricks_decomp *d, size_t step, int stage, int cap, int *npieces, size_t *recv_off,
                             size_t *send_off, size_t *nbytes, int *recv_peer, int *send_peer) {
  if (!d || !npieces || stage < 0 || stage > 2) return -1;
  Decomp3 &D = *d->dec;
  if (D.rank_map.empty()) {
    bricks::set_error("bricks_decomp_shift_plan: call bricks_decomp_rank_map first");
    return -1;
  }
  return guarded([&] {
    const auto stages = D.shiftPieces();
    const auto &st = stages[(size_t)stage];
    *npieces = (int)st.size();
    if (cap <= 0) return;
    if (!recv_off || !send_off || !nbytes || !recv_peer || !send_peer || cap < (int)st.size())
      throw std::runtime_error("bricks_decomp_shift_plan: arrays too small");
    const size_t sb = step * sizeof(double);
    for (size_t i = 0; i < st.size(); ++i) {
      recv_off[i] = (size_t)st[i].recv_pos * sb;
      send_off[i] = (size_t)st[i].send_pos * sb;
      nbytes[i] = (size_t)st[i].len * sb;
      recv_peer[i] = D.rank_map[st[i].from.set];
      send_peer[i] = D.rank_map[(!st[i].from).set];
    }
  });
}

int bricks_decomp_step_create(bricks_decomp *d, bricks_ctx *ctx, int rank, int nranks, int exchange_kind,
                              size_t step, bricks_step **out, char *what, size_t what_len) {
  if (!d || !ctx || !out || exchange_kind < 0 || exchange_kind > 2) return -1;
  Decomp3 &D = *d->dec;
  if (D.rank_map.empty()) {
    bricks::set_error("bricks_decomp_step_create: call bricks_decomp_rank_map first");
    return -1;
  }
  return guarded([&] {
    D.comm.ctx = ctx;
    D.comm.rank = rank;
    D.comm.size = nranks;
    std::string text;
    *out = D.makeStep((Decomp3::StepExchange)exchange_kind, step, &text);
    if (what && what_len) snprintf(what, what_len, "%s", text.c_str());
  });
}

int bricks_decomp4d_create(const unsigned dims[4], unsigned depth, int page_align, bricks_decomp4d **out) {
  if (!dims || !out || !depth || depth % 4) {
    bricks::set_error("bricks_decomp4d_create: bad argument (extents and depth must be multiples of 4)");
    return -1;
  }
  for (int a = 0; a < 4; ++a)
    if (dims[a] % 4 || dims[a] < 2 * depth) {
      bricks::set_error("bricks_decomp4d_create: extent %u is not a multiple of 4 or smaller than two ghost depths",
                        dims[a]);
      return -1;
    }
  return guarded([&] {
    std::unique_ptr<Decomp4> d(new Decomp4({dims[0], dims[1], dims[2], dims[3]}, depth));
    d->page_align = page_align != 0;
    std::vector<BitSet> skin; /* weak/main-4d.cpp:103-106 */
    allneighbors(0, 1, 4, skin);
    skin.erase(skin.begin() + (skin.size() / 2));
    d->initialize(skin);
    *out = new bricks_decomp4d{d.release()};
  });
}

int bricks_decomp4d_destroy(bricks_decomp4d *d) {
  if (d) {
    delete d->dec;
    delete d;
  }
  return 0;
}

int bricks_decomp4d_sizes(const bricks_decomp4d *d, unsigned *nbricks, unsigned sep_pos[3], unsigned gdims[4],
                          int *nregions, int *nskin) {
  if (!d) return -1;
  if (nbricks) *nbricks = d->dec->sep_pos[2];
  for (int a = 0; a < 3; ++a)
    if (sep_pos) sep_pos[a] = d->dec->sep_pos[a];
  for (int a = 0; a < 4; ++a)
    if (gdims) gdims[a] = d->dec->gridDim(a);
  if (nregions) *nregions = (int)d->dec->ghost.size();
  if (nskin) *nskin = (int)d->dec->skinlist.size();
  return 0;
}

int bricks_decomp4d_tables(const bricks_decomp4d *d, unsigned *grid, unsigned *adj, bricks_region *ghost,
                           bricks_region *skin, long *skin_size) {
  if (!d) return -1;
  Decomp4 &D = *d->dec;
  const size_t cells = (size_t)D.gridDim(0) * D.gridDim(1) * D.gridDim(2) * D.gridDim(3);
  if (grid) memcpy(grid, D.gridData(), cells * sizeof(unsigned));
  if (adj) memcpy(adj, D.getBrickInfo().adj, (size_t)D.sep_pos[2] * 81 * sizeof(unsigned));
  if (ghost) flatten(D.ghost, ghost);
  if (skin) flatten(D.skin, skin);
  if (skin_size)
    for (size_t i = 0; i < D.skin_size.size(); ++i) skin_size[i] = D.skin_size[i];
  return 0;
}

int bricks_decomp4d_rank_map(bricks_decomp4d *d, const int procdims[4], const int coo_in[4], uint64_t keys[81],
                             int ranks[81]) {
  if (!d || !procdims || !coo_in) return -1;
  BricksComm c;
  c.ndims = 4;
  c.size = 1;
  for (int a = 0; a < 4; ++a) {
    if (procdims[a] < 1) return -1;
    c.dims[a] = procdims[a];
    c.size *= procdims[a];
  }
  int coo[4] = {coo_in[0], coo_in[1], coo_in[2], coo_in[3]};
  c.rank = cart_rank(c, coo);
  c.ctx = d->dec->comm.ctx;
  d->dec->comm = c;
  d->dec->rank_map.clear();
  populate(c, *d->dec, 0, 1, coo);
  if (keys && ranks) {
    std::vector<BitSet> nb;
    allneighbors(0, 1, 4, nb);
    for (size_t i = 0; i < nb.size(); ++i) {
      keys[i] = nb[i].set;
      ranks[i] = d->dec->rank_map[nb[i].set];
    }
  }
  return 0;
}

int bricks_decomp4d_exchange_plan(bricks_decomp4d *d, size_t step, size_t *recv_off, size_t *send_off, size_t *nbytes,
                                  int *recv_peer, int *send_peer) {
  if (!d || !recv_off || !send_off || !nbytes || !recv_peer || !send_peer) return -1;
  Decomp4 &D = *d->dec;
  if (D.rank_map.empty()) {
    bricks::set_error("bricks_decomp4d_exchange_plan: call bricks_decomp4d_rank_map first");
    return -1;
  }
  const size_t sb = step * sizeof(double);
  for (size_t i = 0; i < D.ghost.size(); ++i) {
    const auto &g = D.ghost[i];
    const auto &s = D.skin[i];
    recv_off[i] = (size_t)(g.pos + g.first_pad) * sb;
    send_off[i] = (size_t)(s.pos + s.first_pad) * sb;
    nbytes[i] = (size_t)(g.len - g.first_pad - g.last_pad) * sb;
    recv_peer[i] = D.rank_map[g.neighbor.set];
    send_peer[i] = D.rank_map[s.neighbor.set];
  }
  return 0;
}

int bricks_copy_brick4(int to_brick, const long dimlist[4], const long padding[4], const long ghost[4], double *arr,
                       const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj, unsigned nbricks) {
  if (!dimlist || !padding || !ghost || !arr || !grid || !brick_dat || !adj) return -1;
  return guarded([&] {
    BrickInfo<4> info(1);
    void *own = info.adj;
    info.adj = (BrickInfo<4>::adjlist) const_cast<unsigned *>(adj);
    info.nbricks = nbricks;
    BrickStorage st;
    st.dat = std::shared_ptr<bElem>(brick_dat, [](bElem *) {});
    st.chunks = nbricks;
    st.step = step;
    Brick4 b(&info, st, 0);
    std::vector<long> dl(dimlist, dimlist + 4), pd(padding, padding + 4), gh(ghost, ghost + 4);
    if (to_brick)
      copyToBrick<4>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b);
    else
      copyFromBrick<4>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b);
    free(own);
  });
}

int bricks_init_grid3(const long dimlist[3], unsigned *grid, unsigned *adj) {
  if (!dimlist || !grid || !adj) return -1;
  return guarded([&] {
    unsigned *g = nullptr;
    BrickInfo<3> info = init_grid<3>(g, {dimlist[0], dimlist[1], dimlist[2]});
    memcpy(grid, g, (size_t)info.nbricks * sizeof(unsigned));
    memcpy(adj, info.adj, (size_t)info.nbricks * 27 * sizeof(unsigned));
    free(g);
    free(info.adj);
  });
}

int bricks_copy_brick3(int to_brick, const long dimlist[3], const long padding[3], const long ghost[3], double *arr,
                       const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj, unsigned nbricks) {
  if (!dimlist || !padding || !ghost || !arr || !grid || !brick_dat || !adj) return -1;
  return guarded([&] {
    View v(adj, nbricks, brick_dat, step);
    Brick3 b(&v.info, v.st, 0);
    std::vector<long> dl(dimlist, dimlist + 3), pd(padding, padding + 3), gh(ghost, ghost + 3);
    if (to_brick)
      copyToBrick<3>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b);
    else
      copyFromBrick<3>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b);
  });
}

int bricks_compare_brick3(const long dimlist[3], const long padding[3], const long ghost[3], double *arr,
                          const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj, unsigned nbricks,
                          int *match, double *max_rel) {
  if (!dimlist || !padding || !ghost || !arr || !grid || !brick_dat || !adj) return -1;
  return guarded([&] {
    View v(adj, nbricks, brick_dat, step);
    Brick3 b(&v.info, v.st, 0);
    std::vector<long> dl(dimlist, dimlist + 3), pd(padding, padding + 3), gh(ghost, ghost + 3);
    if (match) *match = compareBrick<3>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b) ? 1 : 0;
    if (max_rel) *max_rel = maxRelErrBrick<3>(dl, pd, gh, arr, const_cast<unsigned *>(grid), b);
  });
}

int bricks_strong_plan_create(unsigned dom_size, unsigned sdom_size, int nranks, int rank, const bricks_decomp *d,
                              size_t step, bricks_strong_plan **out) {
  if (!d || !out || nranks < 1 || rank < 0 || rank >= nranks || !sdom_size) return -1;
  return guarded([&] {
    const Decomp3 &D = *d->dec;
    const size_t sb = step * sizeof(double);
    std::vector<StrongRegion> regs(D.ghost.size());
    for (size_t i = 0; i < regs.size(); ++i) {
      regs[i].ghost_neighbor = D.ghost[i].neighbor.set;
      regs[i].ghost_off = (size_t)(D.ghost[i].pos + D.ghost[i].first_pad) * sb;
      regs[i].skin_off = (size_t)(D.skin[i].pos + D.skin[i].first_pad) * sb;
      regs[i].len = (size_t)(D.ghost[i].len - D.ghost[i].first_pad - D.ghost[i].last_pad) * sb;
    }
    *out = new bricks_strong_plan{new StrongPlan(dom_size, sdom_size, nranks, rank, regs)};
  });
}

int bricks_strong_step_create(const bricks_strong_plan *p, bricks_ctx *ctx, int rank, const bricks_decomp *d,
                              void *a_base, size_t sub_bytes, bricks_step **out, size_t *sent_bytes) {
  if (!p || !ctx || !d || !a_base || !out) return -1;
  return guarded([&] {
    Decomp3 &D = *d->dec;
    BrickInfo<3> info = D.getBrickInfo();
    const unsigned g[3] = {D.gridDim(0), D.gridDim(1), D.gridDim(2)};
    *out = makeStrongStep(ctx, rank, *p->plan, D.gridData(), g, &info.adj[0][0], info.nbricks, D.ghostBricks(),
                          (char *)a_base, sub_bytes, sent_bytes);
  });
}

int bricks_strong_plan_destroy(bricks_strong_plan *p) {
  if (p) {
    delete p->plan;
    delete p;
  }
  return 0;
}

int bricks_strong_plan_sizes(const bricks_strong_plan *p, unsigned long *first_id, unsigned long *last_id,
                             size_t *nlinks, int *nsend_peers, int *nrecv_peers) {
  if (!p) return -1;
  if (first_id) *first_id = p->plan->first_id;
  if (last_id) *last_id = p->plan->last_id;
  if (nlinks) *nlinks = p->plan->links.size();
  if (nsend_peers) *nsend_peers = (int)p->plan->send.size();
  if (nrecv_peers) *nrecv_peers = (int)p->plan->recv.size();
  return 0;
}

int bricks_strong_plan_links(const bricks_strong_plan *p, bricks_strong_link *out) {
  if (!p || !out) return -1;
  for (size_t i = 0; i < p->plan->links.size(); ++i) {
    const StrongLink &l = p->plan->links[i];
    out[i] = bricks_strong_link{l.src_sub, l.dst_sub, l.src_off, l.dst_off, l.len};
  }
  return 0;
}

static const std::vector<StrongPeer> *peer_list(const bricks_strong_plan *p, int dir) {
  return dir == 0 ? &p->plan->send : (dir == 1 ? &p->plan->recv : nullptr);
}

int bricks_strong_plan_peer(const bricks_strong_plan *p, int dir, int idx, int *rank, size_t *bytes, int *npieces) {
  if (!p || !peer_list(p, dir) || idx < 0 || idx >= (int)peer_list(p, dir)->size()) return -1;
  const StrongPeer &q = (*peer_list(p, dir))[idx];
  if (rank) *rank = q.rank;
  if (bytes) *bytes = q.bytes;
  if (npieces) *npieces = (int)q.pieces.size();
  return 0;
}

int bricks_strong_plan_pieces(const bricks_strong_plan *p, int dir, int idx, bricks_strong_piece *out) {
  if (!p || !out || !peer_list(p, dir) || idx < 0 || idx >= (int)peer_list(p, dir)->size()) return -1;
  const StrongPeer &q = (*peer_list(p, dir))[idx];
  for (size_t i = 0; i < q.pieces.size(); ++i) {
    const StrongPiece &c = q.pieces[i];
    out[i] = bricks_strong_piece{c.sub, c.region, c.off, c.len, c.buf_off, c.sender_id};
  }
  return 0;
}

int bricks_strong_owner(const bricks_strong_plan *p, unsigned long id, int *rank) {
  if (!p || !rank) return -1;
  *rank = p->plan->owner(id);
  return 0;
}

int bricks_strong_neighbor(const bricks_strong_plan *p, unsigned long id, uint64_t set, unsigned long *nid) {
  if (!p || !nid) return -1;
  *nid = p->plan->neighbor_id(id, BitSet(set));
  return 0;
}

int bricks_strong_coords(const bricks_strong_plan *p, unsigned long id, unsigned coords[3]) {
  if (!p || !coords) return -1;
  p->plan->coords(id, coords);
  return 0;
}

int bricks_cart_dims(int nranks, int ndims, int *dims) {
  if (nranks < 1 || ndims < 1 || ndims > 4 || !dims) return -1;
  cart_dims_create(nranks, ndims, dims);
  return 0;
}

int bricks_cart_rank(int ndims, const int *dims, const int *coo, int *rank) {
  if (ndims < 1 || ndims > 4 || !dims || !coo || !rank) return -1;
  BricksComm c;
  c.ndims = ndims;
  for (int a = 0; a < ndims; ++a) c.dims[a] = dims[a];
  *rank = cart_rank(c, coo);
  return 0;
}

int bricks_cart_coords(int ndims, const int *dims, int rank, int *coo) {
  if (ndims < 1 || ndims > 4 || !dims || !coo) return -1;
  BricksComm c;
  c.ndims = ndims;
  for (int a = 0; a < ndims; ++a) c.dims[a] = dims[a];
  cart_coords(c, rank, coo);
  return 0;
}

unsigned long bricks_zmort3(unsigned long c0, unsigned long c1, unsigned long c2) { return ZMORT()[c0][c1][c2].id; }
unsigned long bricks_zmort3_coord(unsigned long id, unsigned long d) { return ZMORT(id, 3)(d); }
unsigned long bricks_zmort3_set(unsigned long id, unsigned long d, unsigned long p) { return ZMORT(id, 3).set(d, p).id; }

int bricks_uniform_fill(double *out, size_t n, uint64_t seed, uint64_t start) {
  if (!out) return -1;
#pragma omp parallel for
  for (long i = 0; i < (long)n; ++i) out[i] = bricks_detail::uniform01(seed, start + (uint64_t)i);
  return 0;
}

}  // extern "C"
