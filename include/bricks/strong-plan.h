// strong-plan.h -- two-level decomposition for strong scaling: a d^3 domain cut into (d/s)^3
// fixed-size subdomains named by Z-Morton ids, contiguous id ranges owned by each rank, and the
// exchange lists that connect them.
//
// Host logic of the reference's strong drivers, restated as a reusable planner:
//   ownership split          strong/args.cpp:104-113  (remainder subdomains go to the low ranks)
//   getrank()                strong/args.cpp:36-56    (periodic neighbour in Morton space -> rank, local index)
//   intra-rank RegionLink    strong/main.cu:201-208   (neighbour's skin[i] -> my ghost[i], device to device)
//   per-peer packed messages strong/main.cu:209-339   (one buffer per peer rank; both sides order the
//                                                      pieces by (sender Morton id, region index))
// All subdomains share ONE BrickDecomp / BrickInfo (strong/main.cu:131-134), so a region is the
// same brick range in every subdomain's storage.  The reference additionally merges overlapping
// skin ranges that go to the same peer; here every (sender, region) pair is its own piece, which
// moves the same bytes into the same ghost bricks with simpler bookkeeping.
#ifndef BRICKS_B200_STRONG_PLAN_H
#define BRICKS_B200_STRONG_PLAN_H

#include <algorithm>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../bricks_b200.h"
#include "bitset.h"
#include "zmort.h"

/// a device-to-device copy between two local subdomain storages (byte offsets inside a storage)
struct StrongLink {
  unsigned src_sub, dst_sub;
  size_t src_off, dst_off, len;
};
/// one piece of a packed per-peer message
struct StrongPiece {
  unsigned sub;    ///< local subdomain index
  size_t off;      ///< byte offset of the region inside that subdomain's storage
  size_t len;      ///< bytes
  size_t buf_off;  ///< byte offset inside the packed buffer
  unsigned long sender_id;
  unsigned region;
};
struct StrongPeer {
  int rank = 0;
  size_t bytes = 0;
  std::vector<StrongPiece> pieces;
};

struct StrongRegion {  ///< what the planner needs from BrickDecomp::ghost[i] / skin[i]
  uint64_t ghost_neighbor;
  size_t ghost_off, skin_off, len;  // bytes
};

class StrongPlan {
 public:
  unsigned subdim;               ///< subdomains per axis
  unsigned long first_id = 0, last_id = 0;  ///< owned Morton ids [first_id, last_id)
  std::vector<StrongLink> links;
  std::vector<StrongPeer> send, recv;  ///< ascending peer rank

  /// regions: one entry per BrickDecomp region i, in order
  StrongPlan(unsigned dom_size, unsigned sdom_size, int nranks, int rank, const std::vector<StrongRegion> &regions)
      : subdim(dom_size / sdom_size), nranks_(nranks), rank_(rank) {
    if (sdom_size == 0 || dom_size % sdom_size || (subdim & (subdim - 1)))
      throw std::runtime_error("StrongPlan: the number of subdomains per axis must be a power of two");
    total_ = (unsigned long)subdim * subdim * subdim;
    if ((unsigned long)nranks > total_) throw std::runtime_error("StrongPlan: more ranks than subdomains");
    range_of(rank, first_id, last_id);

    std::map<int, std::vector<StrongPiece>> sends, recvs;
    for (unsigned long id = first_id; id < last_id; ++id) {
      const unsigned me = (unsigned)(id - first_id);
      for (unsigned i = 0; i < regions.size(); ++i) {
        const StrongRegion &r = regions[i];
        // ghost[i] of `id` is filled from skin[i] of the neighbour in direction ghost[i].neighbor
        const unsigned long from = neighbor_id(id, BitSet(r.ghost_neighbor));
        const int from_rank = owner(from);
        if (from_rank == rank_) {
          links.push_back(StrongLink{(unsigned)(from - first_id), me, r.skin_off, r.ghost_off, r.len});
        } else {
          recvs[from_rank].push_back(StrongPiece{me, r.ghost_off, r.len, 0, from, i});
        }
        // skin[i] of `id` goes to the neighbour in the mirrored direction
        const unsigned long to = neighbor_id(id, !BitSet(r.ghost_neighbor));
        const int to_rank = owner(to);
        if (to_rank != rank_) sends[to_rank].push_back(StrongPiece{me, r.skin_off, r.len, 0, id, i});
      }
    }
    finish(sends, send);
    finish(recvs, recv);
  }

  /// contiguous split of the Morton ids; ranks below total % nranks own one extra
  void range_of(int rank, unsigned long &lo, unsigned long &hi) const {
    const unsigned long base = total_ / nranks_, extra = total_ % nranks_;
    lo = (unsigned long)rank * base + std::min<unsigned long>(rank, extra);
    hi = lo + base + ((unsigned long)rank < extra ? 1 : 0);
  }
  int owner(unsigned long id) const {
    const unsigned long base = total_ / nranks_, extra = total_ % nranks_;
    const unsigned long split = extra * (base + 1);
    return id < split ? (int)(id / (base + 1)) : (int)(extra + (id - split) / base);
  }
  /// Morton id of the periodic neighbour: axis d+1 of the set moves Morton coordinate d
  unsigned long neighbor_id(unsigned long id, BitSet dir) const {
    ZMORT z(id, 3);
    for (int d = 0; d < 3; ++d) {
      long step = dir.get(d + 1) ? 1 : (dir.get(-1 - d) ? -1 : 0);
      if (step) z = z.set(d, (z(d) + subdim + step) % subdim);
    }
    return z.id;
  }
  /// grid coordinate (axis 0 = i) of a subdomain
  void coords(unsigned long id, unsigned c[3]) const {
    ZMORT z(id, 3);
    for (int d = 0; d < 3; ++d) c[d] = (unsigned)z(d);
  }

 private:
  int nranks_, rank_;
  unsigned long total_;

  static void finish(std::map<int, std::vector<StrongPiece>> &by_rank, std::vector<StrongPeer> &out) {
    for (auto &kv : by_rank) {
      StrongPeer p;
      p.rank = kv.first;
      std::sort(kv.second.begin(), kv.second.end(), [](const StrongPiece &a, const StrongPiece &b) {
        return a.sender_id < b.sender_id || (a.sender_id == b.sender_id && a.region < b.region);
      });
      for (StrongPiece &pc : kv.second) {
        pc.buf_off = p.bytes;
        p.bytes += pc.len;
      }
      p.pieces = std::move(kv.second);
      out.push_back(std::move(p));
    }
  }
};

/// The library's sweep schedule (bricks_step_*, csrc/step.cu) for the local subdomains of a
/// StrongPlan: the box plans over the ONE grid / BrickInfo every subdomain shares, and the
/// communication program of strong/main.cu:397-447 -- pack the pieces every peer needs into one
/// buffer per peer (an arena owned by the step), one send/recv per peer, unpack, then the
/// intra-GPU links neighbour's skin[i] -> my ghost[i] (main.cu:201-208).  Subdomain s of the input
/// field lives at a_base + s * sub_bytes; the link tables hold absolute device addresses, so a
/// step always exchanges THAT field (ST_ITER is even: the result ends there).
/// Used by drivers/strong.cpp and, through bricks_strong_step_create, by bench.py and the tests.
inline bricks_step *makeStrongStep(bricks_ctx *ctx, int rank, const StrongPlan &sp, const unsigned *grid,
                                   const unsigned gdims[3], const unsigned *adj, unsigned nbricks,
                                   unsigned ghost_bricks, char *a_base, size_t sub_bytes, size_t *sent_bytes = nullptr) {
  auto require = [](int rc, const char *what) {
    if (rc != 0) throw std::runtime_error(std::string(what) + ": " + bricks_last_error());
  };
  bricks_step *st = nullptr;
  require(bricks_step_create(ctx, grid, gdims, adj, nbricks, ghost_bricks, &st), "bricks_step_create");
  try {
    std::vector<int> peers;
    for (auto &p : sp.send) peers.push_back(p.rank);
    for (auto &p : sp.recv) peers.push_back(p.rank);
    std::sort(peers.begin(), peers.end());
    peers.erase(std::unique(peers.begin(), peers.end()), peers.end());
    auto find = [](const std::vector<StrongPeer> &v, int r) -> const StrongPeer * {
      for (auto &p : v)
        if (p.rank == r) return &p;
      return nullptr;
    };
    std::vector<size_t> so, sb, ro, rb;
    size_t off = 0, sent = 0;
    for (int r : peers) {
      const StrongPeer *p = find(sp.send, r);
      so.push_back(off), sb.push_back(p ? p->bytes : 0);
      off += (sb.back() + 255) / 256 * 256;
      sent += sb.back();
    }
    for (int r : peers) {
      const StrongPeer *p = find(sp.recv, r);
      ro.push_back(off), rb.push_back(p ? p->bytes : 0);
      off += (rb.back() + 255) / 256 * 256;
    }
    if (sent_bytes) *sent_bytes = sent;
    auto link = [](size_t from, size_t to, size_t len) {
      bricks_region_link l;
      l.from = (double *)from;
      l.to = (double *)to;
      l.len = len;
      return l;
    };
    const size_t a0 = (size_t)a_base;
    std::vector<bricks_region_link> local, pack, unpack;
    for (auto &l : sp.links)
      local.push_back(link(a0 + l.src_sub * sub_bytes + l.src_off, a0 + l.dst_sub * sub_bytes + l.dst_off, l.len));
    for (size_t k = 0; k < peers.size(); ++k) {
      if (const StrongPeer *p = find(sp.send, peers[k]))
        for (auto &pc : p->pieces) pack.push_back(link(a0 + pc.sub * sub_bytes + pc.off, so[k] + pc.buf_off, pc.len));
      if (const StrongPeer *p = find(sp.recv, peers[k]))
        for (auto &pc : p->pieces) unpack.push_back(link(ro[k] + pc.buf_off, a0 + pc.sub * sub_bytes + pc.off, pc.len));
    }
    if (!peers.empty()) {
      void *arena = nullptr;
      require(bricks_step_arena(st, off, &arena), "bricks_step_arena");
      require(bricks_step_add_copy(st, pack.data(), (int)pack.size(), BRICKS_BASE_ABS, BRICKS_BASE_ARENA),
              "bricks_step_add_copy");
      bricks_exchange *ex = nullptr;
      require(bricks_exchange_create_v(ctx, rank, (int)peers.size(), ro.data(), rb.data(), peers.data(), so.data(),
                                       sb.data(), peers.data(), &ex), "bricks_exchange_create_v");
      require(bricks_step_add_exchange(st, ex, BRICKS_BASE_ARENA, 1), "bricks_step_add_exchange");
      require(bricks_step_add_copy(st, unpack.data(), (int)unpack.size(), BRICKS_BASE_ARENA, BRICKS_BASE_ABS),
              "bricks_step_add_copy");
    }
    require(bricks_step_add_copy(st, local.data(), (int)local.size(), BRICKS_BASE_ABS, BRICKS_BASE_ABS),
            "bricks_step_add_copy");
  } catch (...) {
    bricks_step_destroy(st);
    throw;
  }
  return st;
}

#endif  // BRICKS_B200_STRONG_PLAN_H
