// brick-decomp.h -- communication-friendly brick numbering of one subdomain and its halo
// exchange on the device.
//
// Drop-in for the reference's include/brick-mpi.h + src/brick-mpi.cpp:
//   BrickDecomp<dim, BDims...>   (brick-mpi.h:178-713)  same public members: ghost, skin,
//       sep_pos, skinlist, skin_size, comm, rank_map, initialize(), exchange(), operator[],
//       getBrickInfo(), exchangeView()
//   ExchangeView                 (brick-mpi.h:82-124)   one logical message per neighbour
//   populate()                   (brick-mpi.h:730-753)  neighbour set -> rank
//   allneighbors(), skin3d_good / skin3d_normal / skin3d_bad   (brick-mpi.cpp:9-136)
//   mpi_stats / mpi_statistics   (brick-mpi.h:758-793)
//   packtime, calltime, waittime, movetime, calctime           (brick-mpi.cpp:7)
// What changed underneath: there is no MPI.  `MPI_Comm` becomes BricksComm, a periodic
// Cartesian process grid bound to one device context of libbricks_b200 (include/
// bricks_b200.h); exchange() hands DEVICE storages to bricks_exchange_run (NCCL send/recv
// over NVLink, or on-device copies when the neighbour is this rank itself).  The numbering
// is bit-exact with the reference (pinned by tests/golden/decomp.json, which was dumped from
// the compiled reference).  The numbering code is new: every region is an axis-aligned box
// of the brick grid and is numbered by one flat odometer loop.
#ifndef BRICKS_B200_BRICK_DECOMP_H
#define BRICKS_B200_BRICK_DECOMP_H

#include <unistd.h>

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstring>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../bricks_b200.h"
#include "bitset.h"
#include "brick.h"

// phase timers of the reference drivers (seconds, accumulated)
inline double packtime = 0, calltime = 0, waittime = 0, movetime = 0, calctime = 0;

/// Replacement for the Cartesian MPI communicator of weak/args.cpp:105-115: a periodic
/// process grid (index 0 slowest, ranks row-major as MPI_Cart_rank numbers them) plus the
/// device context this rank drives.
struct BricksComm {
  int ndims = 0;
  int dims[4] = {1, 1, 1, 1};
  int rank = 0;
  int size = 1;
  bricks_ctx *ctx = nullptr;
};

namespace bricks_detail {
inline void require(int rc, const char *what) {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + bricks_last_error());
}
}  // namespace bricks_detail

/// MPI_Dims_create(nnodes, ndims, dims) with all entries free: balanced factors in
/// non-increasing order (8 -> 2,2,2; 4 -> 2,2,1; 2 -> 2,1,1; 6 -> 3,2,1).
inline void cart_dims_create(int nnodes, int ndims, int *dims) {
  std::vector<int> primes;
  for (int p = 2, n = nnodes; n > 1; ++p)
    while (n % p == 0) {
      primes.push_back(p);
      n /= p;
    }
  std::vector<int> d(ndims, 1);
  for (auto it = primes.rbegin(); it != primes.rend(); ++it) *std::min_element(d.begin(), d.end()) *= *it;
  std::sort(d.begin(), d.end(), [](int a, int b) { return a > b; });
  for (int i = 0; i < ndims; ++i) dims[i] = d[i];
}

/// MPI_Cart_rank on a fully periodic grid
inline int cart_rank(const BricksComm &c, const int *coo) {
  int r = 0;
  for (int a = 0; a < c.ndims; ++a) r = r * c.dims[a] + ((coo[a] % c.dims[a]) + c.dims[a]) % c.dims[a];
  return r;
}

/// MPI_Cart_coords
inline void cart_coords(const BricksComm &c, int rank, int *coo) {
  for (int a = c.ndims - 1; a >= 0; --a) {
    coo[a] = rank % c.dims[a];
    rank /= c.dims[a];
  }
}

inline BricksComm cart_create(int nranks, int rank, int ndims, bricks_ctx *ctx = nullptr) {
  BricksComm c;
  c.ndims = ndims;
  c.size = nranks;
  c.rank = rank;
  c.ctx = ctx;
  cart_dims_create(nranks, ndims, c.dims);
  return c;
}

/// All 3^dim direction sets in the reference's order: per axis "+", "none", "-", axis 1
/// outermost (src/brick-mpi.cpp:9-23).
inline void allneighbors(BitSet cur, long idx, long dim, std::vector<BitSet> &neighbors) {
  long total = 1;
  for (long a = idx; a <= dim; ++a) total *= 3;
  for (long code = 0; code < total; ++code) {
    BitSet s = cur;
    long rem = code;
    for (long a = dim; a >= idx; --a) {  // axis `dim` is the fastest digit
      const long digit = rem % 3;
      rem /= 3;
      if (digit == 0) s.flip(a);
      if (digit == 2) s.flip(-a);
    }
    neighbors.push_back(s);
  }
}

/// Surface orderings (src/brick-mpi.cpp:25-136).  skin3d_good walks the 26 surface regions so
/// that the regions each neighbour needs are contiguous in as few runs as possible.
inline std::vector<BitSet> skin3d_good = {
    {1},       {1, -3},      {1, 2, -3}, {1, 2},      {1, 2, 3},  {2, 3},     {2},        {2, -3},     {-1, 2, -3},
    {-1, 2},   {-1, 2, 3},   {-1, 3},    {-1},        {-3},       {-1, -3},   {-1, -2, -3}, {-1, -2},  {-1, -2, 3},
    {-2, 3},   {-2},         {-2, -3},   {1, -2, -3}, {1, -2},    {1, -2, 3}, {1, 3},     {3}};

namespace bricks_detail {
inline std::vector<BitSet> lexicographic_skin(bool spaced) {
  std::vector<BitSet> out;
  for (int k = -1; k <= 1; ++k)
    for (int j = -1; j <= 1; ++j)
      for (int i = -1; i <= 1; ++i) {
        BitSet s;
        if (i) s.flip(i * 1);
        if (j) s.flip(j * 2);
        if (k) s.flip(k * 3);
        if (spaced && s.set == 0) continue;  // the spaced list has no centre entry of its own
        if (spaced && !out.empty()) out.push_back(BitSet());
        out.push_back(s);
      }
  return out;
}
}  // namespace bricks_detail

inline std::vector<BitSet> skin3d_normal = bricks_detail::lexicographic_skin(false);
inline std::vector<BitSet> skin3d_bad = bricks_detail::lexicographic_skin(true);

/// One logical message per neighbour (brick-mpi.h:82-124).  The reference glues the several
/// storage ranges a neighbour needs into one virtual range with mmap; here the ranges stay
/// where they are and the device exchange walks them (same bytes, same pairing).
struct ExchangeView {
  BricksComm comm;
  std::vector<size_t> seclen;     ///< payload bytes per neighbour
  std::vector<size_t> first_pad;  ///< kept for compatibility (bytes)
  typedef std::vector<std::pair<int, void *>> Dest;
  Dest send, recv;  ///< (peer rank, first range) per neighbour
  bricks_exchange *plan = nullptr;
  bElem *storage_dev = nullptr;

  /// blocking: returns when the ghost bricks of `storage_dev` are up to date
  void exchange() {
    start();
    wait();
  }
  /// enqueue on the communication stream; compute on interior bricks may proceed
  void start() {
    bricks_detail::require(bricks_stream_wait(comm.ctx, BRICKS_STREAM_COMM, BRICKS_STREAM_COMPUTE), "exchange");
    bricks_detail::require(bricks_exchange_run(plan, (double *)storage_dev, BRICKS_STREAM_COMM), "exchange");
  }
  /// make the compute stream wait for the exchange (no host block)
  void wait() { bricks_detail::require(bricks_stream_wait(comm.ctx, BRICKS_STREAM_COMPUTE, BRICKS_STREAM_COMM), "exchange"); }
};

/// SHIFT exchange (brick-mpi.h:126-176, 587-681): one stage per axis, two messages per stage.  What
/// an edge or corner neighbour owns arrives in several hops through the ghost zones of the face
/// neighbours, so a rank talks to 6 peers instead of 26.  The reference glues the pieces of a
/// message into one virtual range with mmap; here a message is a run of (send range, receive range)
/// pairs of equal length that both ends derive from the same tables, issued inside one NCCL group per
/// stage (self-neighbours become device copies).  Stages are ordered on the communication stream: a
/// stage forwards ghost bricks the previous one received.
struct MultiStageExchangeView {
  BricksComm comm;
  std::vector<bricks_exchange *> stages;  ///< owned by the BrickDecomp that made the view
  bElem *storage_dev = nullptr;

  void exchange() {
    start();
    wait();
  }
  void start() {
    bricks_detail::require(bricks_stream_wait(comm.ctx, BRICKS_STREAM_COMM, BRICKS_STREAM_COMPUTE), "exchange");
    for (bricks_exchange *st : stages)
      bricks_detail::require(bricks_exchange_run(st, (double *)storage_dev, BRICKS_STREAM_COMM), "exchange");
  }
  void wait() { bricks_detail::require(bricks_stream_wait(comm.ctx, BRICKS_STREAM_COMPUTE, BRICKS_STREAM_COMM), "exchange"); }
};

/// One (send range, receive range) pair of a SHIFT-exchange stage, in bricks.
struct ShiftPiece {
  unsigned recv_pos, send_pos, len;
  BitSet from;  ///< direction received from; the send goes to the mirrored direction
};

template <unsigned dim, unsigned... BDims>
class BrickDecomp {
 public:
  /// One contiguous run of bricks exchanged with one neighbour
  typedef struct {
    BitSet neighbor;     ///< who it is exchanged with
    unsigned skin_st;    ///< first skinlist entry of the run
    unsigned skin_ed;    ///< one past the last skinlist entry
    unsigned pos;        ///< first brick
    unsigned len;        ///< number of bricks
    unsigned first_pad;  ///< padding bricks at the start / end (page-aligned builds only)
    unsigned last_pad;
  } g_region;
  std::vector<g_region> ghost;  ///< where neighbour data lands
  std::vector<g_region> skin;   ///< what is sent, paired by index with ghost
  unsigned sep_pos[3];          ///< [1,sep0) interior, [sep0,sep1) skin, [sep1,sep2) ghost
  std::vector<BitSet> skinlist;
  std::vector<long> skin_size;

  BricksComm comm;
  std::unordered_map<uint64_t, int> rank_map;  ///< direction set -> rank
  /// Region starts rounded up to whole pages (the reference's default build) or dense numbering
  /// (its -DDECOMP_PAGEUNALIGN build, which the CUDA drivers use: weak/CMakeLists.txt:32).  The two
  /// differ only for bricks smaller than a page (4x4x4x4 fp64 = 2 KiB).  Set before initialize().
#ifdef DECOMP_PAGEUNALIGN
  bool page_align = false;
#else
  bool page_align = true;
#endif

  /// dims: interior extent in elements, contiguous axis first; depth: ghost depth in elements
  BrickDecomp(const std::vector<unsigned> &dims, const unsigned depth, unsigned numfield = 1)
      : numfield(numfield) {
    assert(dims.size() == dim);
    unsigned bd[dim] = {BDims...};
    for (unsigned a = 0; a < dim; ++a) {
      const unsigned b = bd[dim - 1 - a];  // template lists are contiguous axis LAST
      assert(depth % b == 0 && dims[a] % b == 0);
      nb[a] = dims[a] / b;
      gd[a] = depth / b;
    }
  }
  BrickDecomp(const BrickDecomp &) = delete;
  BrickDecomp &operator=(const BrickDecomp &) = delete;
  ~BrickDecomp() {
    for (auto &kv : plans) bricks_exchange_destroy(kv.second);
    for (auto &kv : shift_plans)
      for (bricks_exchange *ex : kv.second) bricks_exchange_destroy(ex);
    delete[] grid;
    if (bInfo) {
      free(bInfo->adj);
      delete bInfo;
    }
  }

  void initialize(const std::vector<BitSet> &skins) {
    calltime = waittime = 0;
    skinlist = skins;
    cells = 1;
    for (unsigned a = 0; a < dim; ++a) {
      stride[a] = cells;
      td[a] = nb[a] + 2 * gd[a];
      cells *= td[a];
    }
    delete[] grid;
    grid = new unsigned[cells];

    const long pagesize = sysconf(_SC_PAGESIZE);
    const long bsize = (long)cal_size<BDims...>::value * (long)sizeof(bElem) * numfield;
    if (std::max(bsize, pagesize) % std::min(bsize, pagesize) != 0)
      throw std::runtime_error("brick size must be a factor/multiple of pagesize.");
    // bricks smaller than a page: the first pagesize / bsize storage slots are reserved (slot 0 is
    // the sentinel) and, in the page-aligned numbering, every region starts on a page boundary
    const long reserved = bsize < pagesize ? pagesize / bsize : 1;
    const long factor = page_align ? reserved : 1;
    unsigned pos = (unsigned)reserved;
    auto round_up = [factor](unsigned p) { return (unsigned)((p + factor - 1) / factor * factor); };
    auto pad_of = [factor, this](BitSet region) -> unsigned {
      return (unsigned)(factor - (region_bricks(region) + factor - 1) % factor - 1);
    };

    // a region is padded at its front when it shares more with its successor than with its
    // predecessor, so that runs sent to one neighbour stay page-contiguous
    const long ns = (long)skinlist.size();
    std::vector<bool> pad_first(ns);
    for (long i = 0; i < ns; ++i) {
      const BitSet prev = i ? skinlist[i - 1] : BitSet();
      const BitSet next = i + 1 < ns ? skinlist[i + 1] : BitSet();
      pad_first[i] = (prev & skinlist[i]).size() < (skinlist[i] & next).size();
    }

    number_box(BitSet(), BitSet(), pos);
    pos = round_up(pos);
    sep_pos[0] = pos;

    std::vector<unsigned> start(ns + 1);
    start[0] = pos;
    skin_size.assign(ns, 0);
    for (long i = 0; i < ns; ++i) {
      const unsigned before = pos;
      if (pad_first[i]) pos += pad_of(skinlist[i]);
      if (skinlist[i].set != 0) {
        number_box(BitSet(), skinlist[i], pos);
        pos = round_up(pos);
      }
      start[i + 1] = pos;
      skin_size[i] = (long)pos - (long)before;
    }
    sep_pos[1] = pos;

    // ghost copies: for neighbour n, every surface region that contains the mirrored
    // direction set, in skinlist order; consecutive entries form one exchanged run
    std::vector<BitSet> nbrs;
    allneighbors(BitSet(), 1, dim, nbrs);
    ghost.clear();
    skin.clear();
    for (const BitSet n : nbrs) {
      if (n.set == 0) continue;
      const BitSet mirrored = !n;
      long open = -1;
      g_region g{}, s{};
      for (long l = 0; l <= ns; ++l) {
        const bool member = l < ns && mirrored <= skinlist[l];
        if (member) {
          if (open < 0) {
            open = l;
            g.neighbor = n;
            s.neighbor = mirrored;
            g.skin_st = s.skin_st = (unsigned)l;
            g.first_pad = s.first_pad = pad_first[l] ? pad_of(skinlist[l]) : 0;
            g.pos = pos;
            s.pos = start[l];
          }
          if (pad_first[l]) pos += pad_of(skinlist[l]);
          number_box(n, skinlist[l], pos);
          pos = round_up(pos);
        } else if (open >= 0) {
          g.last_pad = s.last_pad = pad_first[l - 1] ? 0 : pad_of(skinlist[l - 1]);
          g.skin_ed = s.skin_ed = (unsigned)l;
          g.len = pos - g.pos;
          s.len = start[l] - s.pos;
          ghost.push_back(g);
          skin.push_back(s);
          open = -1;
        }
      }
    }
    sep_pos[2] = pos;

    if (bInfo == nullptr) bInfo = new BrickInfo<dim>(pos);
    // rows of the sentinel brick and of padding slots stay zero: they have no neighbours
    memset(bInfo->adj[0], 0, sizeof(unsigned) * NSLOT * (size_t)pos);
    for (unsigned c = 0; c < cells; ++c) fill_adjacency(c, bInfo->adj[grid[c]]);
    for (auto &kv : plans) bricks_exchange_destroy(kv.second);
    plans.clear();
    for (auto &kv : shift_plans)
      for (bricks_exchange *ex : kv.second) bricks_exchange_destroy(ex);
    shift_plans.clear();
  }

  /// grid lookup including the ghost band: decomp[k][j][i]
  struct GridRow {
    const BrickDecomp *self;
    unsigned ref, axis;
    inline GridRow operator[](int i) const { return GridRow{self, ref + (unsigned)i * self->stride[axis - 1], axis - 1}; }
    inline operator unsigned() const { return self->grid[ref]; }
  };
  GridRow operator[](int i) const { return GridRow{this, 0, dim}[i]; }

  BrickInfo<dim> getBrickInfo() { return *bInfo; }
  const unsigned *gridData() const { return grid; }
  unsigned gridDim(unsigned axis) const { return td[axis]; }
  unsigned ghostBricks() const { return gd[0]; }

  /// Halo exchange of one DEVICE storage created from this decomposition (blocking, like the
  /// reference's BrickDecomp::exchange(BrickStorage&), brick-mpi.h:466-495).
  void exchange(BrickStorage &storage_dev) {
    ExchangeView v = exchangeView(storage_dev);
    v.exchange();
    bricks_detail::require(bricks_sync(comm.ctx, BRICKS_STREAM_COMPUTE), "exchange");
  }

  /// Per-region exchange plan bound to a device storage; exchange() / start() / wait()
  ExchangeView exchangeView(BrickStorage storage_dev) {
    if (!comm.ctx) throw std::runtime_error("BrickDecomp::comm has no device context");
    const size_t step_bytes = storage_dev.step * sizeof(bElem);
    ExchangeView v;
    v.comm = comm;
    v.storage_dev = storage_dev.dat.get();
    std::vector<BitSet> nbrs;
    allneighbors(BitSet(), 1, dim, nbrs);
    for (const BitSet n : nbrs) {
      if (n.set == 0) continue;
      size_t len = 0, fp = 0;
      void *first_recv = nullptr, *first_send = nullptr;
      for (size_t i = 0; i < ghost.size(); ++i)
        if (ghost[i].neighbor.set == n.set) {
          if (!first_recv) {
            fp = ghost[i].first_pad * step_bytes;
            first_recv = (char *)v.storage_dev + ghost[i].pos * step_bytes;
            first_send = (char *)v.storage_dev + skin[i].pos * step_bytes;
          }
          len += (size_t)(ghost[i].len - ghost[i].first_pad - ghost[i].last_pad) * step_bytes;
        }
      v.seclen.push_back(len);
      v.first_pad.push_back(fp);
      v.recv.push_back(std::make_pair(rank_map[n.set], first_recv));
      v.send.push_back(std::make_pair(rank_map[(!n).set], first_send));
    }
    v.plan = plan_for(storage_dev.step);
    return v;
  }

  /// The SHIFT exchange as piece tables: result[d-1] lists the pieces of stage d (axis d), grouped
  /// by direction in allneighbors() order.  Per direction n the receive side is the ghost regions n
  /// owns followed by every ghost region not yet filled whose owner is n plus axes already
  /// exchanged; the send side is the skin facing the mirrored direction followed, for those further
  /// regions, by the same skin sections inside this rank's ghost copy of the (owner minus n)
  /// neighbour (brick-mpi.h:611-664).  Both lists cover the same bytes in the same order; they are
  /// cut at each other's boundaries so that every piece is one contiguous range on both ends.
  std::vector<std::vector<ShiftPiece>> shiftPieces() const {
    typedef std::pair<unsigned, unsigned> Run;  // first brick, count
    std::vector<std::vector<ShiftPiece>> out;
    std::vector<BitSet> nbrs;
    allneighbors(BitSet(), 1, dim, nbrs);
    BitSet done;
    for (long d = 1; d <= (long)dim; ++d) {
      BitSet axis;
      axis.flip(d);
      axis.flip(-d);
      out.emplace_back();
      for (const BitSet n : nbrs) {
        if (n.set == 0 || !(n <= axis)) continue;
        std::vector<Run> recv, send;
        auto push = [](std::vector<Run> &v, unsigned pos, unsigned len) {
          if (!len) return;
          if (!v.empty() && v.back().first + v.back().second == pos)
            v.back().second += len;  // contiguous with the previous run
          else
            v.push_back(Run(pos, len));
        };
        const BitSet mirrored = !n;
        for (const auto &g : ghost)
          if (g.neighbor.set == n.set) push(recv, g.pos + g.first_pad, g.len - g.first_pad - g.last_pad);
        for (const auto &k : skin)
          if (k.neighbor.set == mirrored.set) push(send, k.pos + k.first_pad, k.len - k.first_pad - k.last_pad);
        const BitSet reach = n | done;
        for (const auto &g : ghost) {
          if (g.neighbor.set == n.set || !(g.neighbor <= reach) || g.neighbor <= done) continue;
          if (g.first_pad || g.last_pad) throw std::runtime_error("shiftPieces: padded regions are not supported");
          push(recv, g.pos, g.len);
          const BitSet via = g.neighbor ^ n;
          for (unsigned sec = g.skin_st; sec < g.skin_ed; ++sec)
            for (const auto &h : ghost)
              if (h.neighbor.set == via.set && sec >= h.skin_st && sec < h.skin_ed) {
                unsigned pos = h.pos;
                for (unsigned j = h.skin_st; j < sec; ++j) pos += (unsigned)skin_size[j];
                push(send, pos, (unsigned)skin_size[sec]);
              }
        }
        // cut both run lists at each other's boundaries
        size_t ir = 0, is = 0;
        unsigned ur = 0, us = 0;  // bricks of the current runs already paired
        while (ir < recv.size() && is < send.size()) {
          const unsigned l = std::min(recv[ir].second - ur, send[is].second - us);
          ShiftPiece p;
          p.recv_pos = recv[ir].first + ur;
          p.send_pos = send[is].first + us;
          p.len = l;
          p.from = n;
          out.back().push_back(p);
          ur += l;
          us += l;
          if (ur == recv[ir].second) ++ir, ur = 0;
          if (us == send[is].second) ++is, us = 0;
        }
        if (ir != recv.size() || is != send.size())
          throw std::runtime_error("shiftPieces: send and receive sides of a stage differ in size");
      }
      done = done | axis;
    }
    return out;
  }

  /// How the halo exchange of a step travels (makeStep).
  enum StepExchange {
    STEP_REGIONS = 0,  ///< one send/recv pair per region straight out of the storage (exchange(), :466-495)
    STEP_PACKED = 1,   ///< one message per peer rank: the role of ExchangeView (brick-mpi.h:96-123, 525-580);
                       ///< the mmap views become a pack kernel, one send/recv per peer and an unpack kernel
    STEP_SHIFT = 2     ///< MultiStageExchangeView (brick-mpi.h:587-681): 3 stages, 6 messages
  };

  /// The library's sweep schedule for this decomposition (bricks_step_*, csrc/step.cu): the box plans
  /// of the overlapped schedule over ONE device copy of grid / adj, and the communication program
  /// of the chosen exchange.  Replaces the body of `brick_func` (weak/main.cu:246-287); used by the
  /// C++ drivers and, through bricks_decomp_step_create, by bench.py and the tests.  The caller
  /// owns the result (bricks_step_destroy).  `what`, if given, receives a one-line description.
  bricks_step *makeStep(StepExchange kind, size_t step_elems, std::string *what = nullptr) {
    static_assert(dim == 3, "the sweep kernels are 3-D");
    if (!comm.ctx) throw std::runtime_error("BrickDecomp::comm has no device context");
    using bricks_detail::require;
    bricks_step *st = nullptr;
    require(bricks_step_create(comm.ctx, grid, td, &bInfo->adj[0][0], bInfo->nbricks, gd[0], &st), "bricks_step_create");
    try {
      const size_t sb = step_elems * sizeof(bElem);
      const size_t n = ghost.size();
      std::vector<size_t> roff(n), soff(n), bytes(n);
      std::vector<int> rpeer(n), speer(n);
      for (size_t i = 0; i < n; ++i) {
        roff[i] = (size_t)(ghost[i].pos + ghost[i].first_pad) * sb;
        soff[i] = (size_t)(skin[i].pos + skin[i].first_pad) * sb;
        bytes[i] = (size_t)(ghost[i].len - ghost[i].first_pad - ghost[i].last_pad) * sb;
        rpeer[i] = rank_map.at(ghost[i].neighbor.set);
        speer[i] = rank_map.at(skin[i].neighbor.set);
      }
      char text[160];
      if (kind == STEP_REGIONS) {
        bricks_exchange *ex = nullptr;
        require(bricks_exchange_create(comm.ctx, comm.rank, (int)n, roff.data(), soff.data(), bytes.data(), rpeer.data(),
                                       speer.data(), &ex), "bricks_exchange_create");
        require(bricks_step_add_exchange(st, ex, BRICKS_BASE_FIELD, 1), "bricks_step_add_exchange");
        snprintf(text, sizeof text, "direct (one ncclSend/ncclRecv pair per region, %zu regions)", n);
      } else if (kind == STEP_SHIFT) {
        std::string counts;
        for (const auto &stage : shiftPieces()) {
          const size_t m = stage.size();
          std::vector<size_t> ro(m), so(m), nb_(m);
          std::vector<int> rp(m), sp(m);
          for (size_t i = 0; i < m; ++i) {
            ro[i] = (size_t)stage[i].recv_pos * sb;
            so[i] = (size_t)stage[i].send_pos * sb;
            nb_[i] = (size_t)stage[i].len * sb;
            rp[i] = rank_map.at(stage[i].from.set);
            sp[i] = rank_map.at((!stage[i].from).set);
          }
          bricks_exchange *ex = nullptr;
          require(bricks_exchange_create(comm.ctx, comm.rank, (int)m, ro.data(), so.data(), nb_.data(), rp.data(),
                                         sp.data(), &ex), "bricks_exchange_create");
          require(bricks_step_add_exchange(st, ex, BRICKS_BASE_FIELD, 1), "bricks_step_add_exchange");
          counts += (counts.empty() ? "" : "+") + std::to_string(m);
        }
        snprintf(text, sizeof text, "shift (3 stages, 6 messages, %s pieces)", counts.c_str());
      } else {
        // One message per peer.  Region i travels with MPI tag i in the reference; here the regions
        // of a peer are packed in the order i on both sides (region i of the sender is region i of
        // the receiver: every rank holds the same tables).  Self neighbours stay one copy table.
        std::vector<int> peers;
        for (size_t i = 0; i < n; ++i)
          for (int p : {rpeer[i], speer[i]})
            if (p != comm.rank && std::find(peers.begin(), peers.end(), p) == peers.end()) peers.push_back(p);
        std::sort(peers.begin(), peers.end());
        std::vector<bricks_region_link> pack, unpack, self;
        std::vector<size_t> s_off, s_bytes, r_off, r_bytes;
        size_t off = 0;
        auto link = [](size_t from, size_t to, size_t len) {
          bricks_region_link l;
          l.from = (double *)from;
          l.to = (double *)to;
          l.len = len;
          return l;
        };
        for (int p : peers) {
          s_off.push_back(off);
          for (size_t i = 0; i < n; ++i)
            if (speer[i] == p && bytes[i]) {
              pack.push_back(link(soff[i], off, bytes[i]));
              off += bytes[i];
            }
          s_bytes.push_back(off - s_off.back());
          off = (off + 255) / 256 * 256;
        }
        for (int p : peers) {
          r_off.push_back(off);
          for (size_t i = 0; i < n; ++i)
            if (rpeer[i] == p && bytes[i]) {
              unpack.push_back(link(off, roff[i], bytes[i]));
              off += bytes[i];
            }
          r_bytes.push_back(off - r_off.back());
          off = (off + 255) / 256 * 256;
        }
        for (size_t i = 0; i < n; ++i)
          if (rpeer[i] == comm.rank && bytes[i]) self.push_back(link(soff[i], roff[i], bytes[i]));
        if (!peers.empty()) {
          void *arena = nullptr;
          require(bricks_step_arena(st, off, &arena), "bricks_step_arena");
          require(bricks_step_add_copy(st, pack.data(), (int)pack.size(), BRICKS_BASE_FIELD, BRICKS_BASE_ARENA),
                  "bricks_step_add_copy");
          bricks_exchange *ex = nullptr;
          require(bricks_exchange_create_v(comm.ctx, comm.rank, (int)peers.size(), r_off.data(), r_bytes.data(),
                                           peers.data(), s_off.data(), s_bytes.data(), peers.data(), &ex),
                  "bricks_exchange_create_v");
          require(bricks_step_add_exchange(st, ex, BRICKS_BASE_ARENA, 1), "bricks_step_add_exchange");
          require(bricks_step_add_copy(st, unpack.data(), (int)unpack.size(), BRICKS_BASE_ARENA, BRICKS_BASE_FIELD),
                  "bricks_step_add_copy");
        }
        require(bricks_step_add_copy(st, self.data(), (int)self.size(), BRICKS_BASE_FIELD, BRICKS_BASE_FIELD),
                "bricks_step_add_copy");
        snprintf(text, sizeof text, "packed (one message per peer rank: %zu peers, %zu regions packed, %zu self regions)",
                 peers.size(), pack.size(), self.size());
      }
      if (what) *what = text;
    } catch (...) {
      bricks_step_destroy(st);
      throw;
    }
    return st;
  }

  /// SHIFT exchange bound to a device storage (brick-mpi.h:587: multiStageExchangeView)
  MultiStageExchangeView multiStageExchangeView(BrickStorage storage_dev) {
    if (!comm.ctx) throw std::runtime_error("BrickDecomp::comm has no device context");
    MultiStageExchangeView v;
    v.comm = comm;
    v.storage_dev = storage_dev.dat.get();
    auto it = shift_plans.find(storage_dev.step);
    if (it == shift_plans.end()) {
      const size_t sb = storage_dev.step * sizeof(bElem);
      std::vector<bricks_exchange *> made;
      for (const auto &stage : shiftPieces()) {
        const size_t n = stage.size();
        std::vector<size_t> roff(n), soff(n), bytes(n);
        std::vector<int> rpeer(n), speer(n);
        for (size_t i = 0; i < n; ++i) {
          roff[i] = (size_t)stage[i].recv_pos * sb;
          soff[i] = (size_t)stage[i].send_pos * sb;
          bytes[i] = (size_t)stage[i].len * sb;
          rpeer[i] = rank_map.at(stage[i].from.set);
          speer[i] = rank_map.at((!stage[i].from).set);
        }
        bricks_exchange *ex = nullptr;
        bricks_detail::require(bricks_exchange_create(comm.ctx, comm.rank, (int)n, roff.data(), soff.data(),
                                                      bytes.data(), rpeer.data(), speer.data(), &ex),
                               "bricks_exchange_create");
        made.push_back(ex);
      }
      it = shift_plans.emplace(storage_dev.step, made).first;
    }
    v.stages = it->second;
    return v;
  }

 private:
  static constexpr unsigned NSLOT = static_power<3, dim>::value;
  std::map<size_t, std::vector<bricks_exchange *>> shift_plans;  ///< SHIFT exchange stages by storage step
  unsigned nb[dim];      ///< interior extent in bricks
  unsigned gd[dim];      ///< ghost depth in bricks
  unsigned td[dim];      ///< extent incl. ghost
  unsigned stride[dim];  ///< grid strides in cells
  unsigned cells = 0;
  unsigned *grid = nullptr;
  unsigned numfield;
  BrickInfo<dim> *bInfo = nullptr;
  std::map<size_t, bricks_exchange *> plans;  ///< by storage step

  /// bricks in a surface region: ghost depth along named axes, interior minus both bands else
  long region_bricks(BitSet region) const {
    long n = 1;
    for (unsigned a = 0; a < dim; ++a)
      n *= (region.get(a + 1) || region.get(-(long)(a + 1))) ? (long)gd[a] : (long)nb[a] - 2 * (long)gd[a];
    return n;
  }

  /// Number the bricks of `region` of the subdomain displaced by `owner` (0 = this one),
  /// axis 0 fastest.  Along axis a the region spans [g, 2g) for "-", [n, n+g) for "+" (the
  /// band touching that face; "-" wins if both are set, as in the reference) and [2g, n)
  /// otherwise; an owner displaced by +-1 shifts the box by +-n bricks.
  void number_box(BitSet owner, BitSet region, unsigned &pos) {
    long lo[dim], cnt[dim];
    for (unsigned a = 0; a < dim; ++a) {
      const long ax = a + 1;
      if (region.get(-ax)) {
        lo[a] = gd[a];
        cnt[a] = gd[a];
      } else if (region.get(ax)) {
        lo[a] = nb[a];
        cnt[a] = gd[a];
      } else {
        lo[a] = 2 * (long)gd[a];
        cnt[a] = (long)nb[a] - lo[a];
      }
      if (owner.get(ax)) lo[a] += nb[a];
      if (owner.get(-ax)) lo[a] -= nb[a];
      if (cnt[a] <= 0) return;
    }
    long idx[dim];
    for (unsigned a = 0; a < dim; ++a) idx[a] = 0;
    while (true) {
      long cell = 0;
      for (unsigned a = 0; a < dim; ++a) cell += (lo[a] + idx[a]) * (long)stride[a];
      grid[cell] = pos++;
      unsigned a = 0;
      while (a < dim && ++idx[a] == cnt[a]) idx[a++] = 0;
      if (a == dim) break;
    }
  }

  /// adjacency row of the brick at grid cell c: slot digits are outermost axis first,
  /// neighbours outside the grid resolve to the sentinel brick 0
  void fill_adjacency(unsigned c, unsigned *row) const {
    long coo[dim];
    for (unsigned a = 0; a < dim; ++a) coo[a] = (c / stride[a]) % td[a];
    for (unsigned slot = 0; slot < NSLOT; ++slot) {
      unsigned rem = slot;
      long cell = 0;
      bool inside = true;
      for (unsigned a = 0; a < dim; ++a) {
        const long x = coo[a] + (long)(rem % 3) - 1;
        rem /= 3;
        inside = inside && x >= 0 && x < (long)td[a];
        cell += x * (long)stride[a];
      }
      row[slot] = inside ? grid[cell] : 0;
    }
  }

  bricks_exchange *plan_for(size_t step) {
    auto it = plans.find(step);
    if (it != plans.end()) return it->second;
    const size_t sb = step * sizeof(bElem);
    const size_t n = ghost.size();
    std::vector<size_t> roff(n), soff(n), bytes(n);
    std::vector<int> rpeer(n), speer(n);
    for (size_t i = 0; i < n; ++i) {
      roff[i] = (size_t)(ghost[i].pos + ghost[i].first_pad) * sb;
      soff[i] = (size_t)(skin[i].pos + skin[i].first_pad) * sb;
      bytes[i] = (size_t)(ghost[i].len - ghost[i].first_pad - ghost[i].last_pad) * sb;
      rpeer[i] = rank_map.at(ghost[i].neighbor.set);
      speer[i] = rank_map.at(skin[i].neighbor.set);
    }
    bricks_exchange *ex = nullptr;
    bricks_detail::require(bricks_exchange_create(comm.ctx, comm.rank, (int)n, roff.data(), soff.data(), bytes.data(),
                                                  rpeer.data(), speer.data(), &ex),
                           "bricks_exchange_create");
    plans[step] = ex;
    return ex;
  }
};

/// rank_map[direction set] = rank of that neighbour on the periodic process grid.  Same
/// signature and sign convention as the reference (brick-mpi.h:730-753): axis d of the set
/// moves Cartesian coordinate coo[dim - d], "+d" to coordinate - 1 and "-d" to + 1.
template <unsigned dim, unsigned... BDims>
void populate(BricksComm &comm, BrickDecomp<dim, BDims...> &bDecomp, BitSet neighbor, int d, int *coo) {
  std::vector<BitSet> sets;
  allneighbors(neighbor, d, dim, sets);
  for (const BitSet s : sets) {
    int c[dim];
    for (unsigned a = 0; a < dim; ++a) c[a] = coo[a];
    for (int ax = d; ax <= (int)dim; ++ax) {
      if (s.get(ax)) c[dim - ax] -= 1;
      if (s.get(-ax)) c[dim - ax] += 1;
    }
    bDecomp.rank_map[s.set] = cart_rank(comm, c);
  }
}

typedef struct {
  double min, max, avg, sigma;
} mpi_stats;

/// Reduce one per-rank number to [min, avg, max] (sigma) (brick-mpi.h:768-785).  Ranks of this
/// build live in one process (one host thread per device), so the caller passes all values.
inline mpi_stats mpi_statistics(const std::vector<double> &per_rank) {
  mpi_stats r = {0, 0, 0, 0};
  const size_t n = per_rank.size();
  if (!n) return r;
  r.min = r.max = per_rank[0];
  double sum = 0, sq = 0;
  for (double v : per_rank) {
    r.min = std::min(r.min, v);
    r.max = std::max(r.max, v);
    sum += v;
    sq += v * v;
  }
  r.avg = sum / n;
  r.sigma = std::sqrt(std::max(0.0, (sq - sum * sum / n) / std::max<size_t>(n - 1, 1)));
  return r;
}
inline mpi_stats mpi_statistics(double stats, const BricksComm &) { return mpi_statistics(std::vector<double>{stats}); }

inline std::ostream &operator<<(std::ostream &os, const mpi_stats &s) {
  return os << "[" << s.min << ", " << s.avg << ", " << s.max << "]" << " (σ: " << s.sigma << ")";
}

#endif  // BRICKS_B200_BRICK_DECOMP_H
