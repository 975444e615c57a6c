// Shared internals of libbricks_b200.so (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <nccl.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "../../include/bricks_b200.h"

namespace bricks {

void set_error(const char *fmt, ...);
extern std::atomic<unsigned long long> g_launches; /* host threads drive one device each */

#define BRK_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      bricks::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); \
      return (int)e_;                                                                   \
    }                                                                                   \
  } while (0)

#define BRK_NCCL(expr)                                                                  \
  do {                                                                                  \
    ncclResult_t r_ = (expr);                                                           \
    if (r_ != ncclSuccess) {                                                            \
      bricks::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, ncclGetErrorString(r_)); \
      return 1000 + (int)r_;                                                            \
    }                                                                                   \
  } while (0)

#define BRK_ARG(cond)                                                                   \
  do {                                                                                  \
    if (!(cond)) {                                                                      \
      bricks::set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond);          \
      return -1;                                                                        \
    }                                                                                   \
  } while (0)

constexpr int kNumStreams = 4;
constexpr int kNumTimers = 16;

}  // namespace bricks

struct bricks_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t streams[bricks::kNumStreams] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t xev = nullptr; /* cross-stream dependency */
  cudaEvent_t t0[bricks::kNumTimers] = {}, t1[bricks::kNumTimers] = {};
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;
};

/* One work item of the marching kernels: a column of nk brick layers at grid (gi..gi+1, gj). */
struct MarchItem {
  unsigned short gi, gj, gk0, nk;
  unsigned valid1; /* second brick of the pair is inside the box */
  unsigned pad_;
};

/* kernel parameters of the sweep kernels (__grid_constant__) */
struct FieldDev;
struct MarchParams {
  const MarchItem *items;
  const unsigned *grid;
  const unsigned *adj;
  const FieldDev *fields;
  unsigned gd0, gd1;
  double c[27];
};

/* Tensor maps of one input field for one kernel variant.  main: P-plane pieces; roll: R-plane
 * pieces used for the pre-/post-roll of a column.  [0]=centre 8x8, [1]=i-face HWx8,
 * [2]=j-face 8xR, [3]=corner HWxR.  pair: the centre and j-face boxes over TWO bricks that are
 * consecutive in storage (main stages only): the two pieces of a pair sit back to back in the
 * stage, so one copy fills both. */
struct alignas(64) MapSet {
  /* [4] / pair[2]: the swizzled four-row j-face box of the 25-point star (march_dev.cuh, JSWZ) */
  CUtensorMap main[5];
  CUtensorMap roll[5];
  CUtensorMap pair[3]; /* [0]=centres of bricks b, b+1; [1]=j-face rows of bricks b, b+1; [2]=swizzled j-face rows */
};

struct alignas(64) FieldDev {
  MapSet maps;
  double *out;
  unsigned long long out_step;
  const double *in; /* same field the maps describe: per-lane cp.async of halo pieces */
  unsigned long long in_step;
  unsigned long long pad[4];
};
static_assert(sizeof(FieldDev) % 64 == 0, "tensor maps in a FieldDev array must stay 64-byte aligned");

struct FieldKey {
  int variant;
  std::vector<const void *> ptrs; /* in0,out0,in1,out1,... */
  size_t in_step, out_step;
  bool operator<(const FieldKey &o) const {
    if (variant != o.variant) return variant < o.variant;
    if (in_step != o.in_step) return in_step < o.in_step;
    if (out_step != o.out_step) return out_step < o.out_step;
    return ptrs < o.ptrs;
  }
};

struct bricks_plan {
  bricks_ctx *ctx = nullptr;
  unsigned gdims[3] = {0, 0, 0};
  unsigned nbricks = 0;
  unsigned *grid_dev = nullptr;
  unsigned *adj_dev = nullptr;
  /* a derived plan (bricks_plan_derive) borrows the tables of its parent: one device copy of
   * grid / adj per subdomain, however many box plans its schedule needs */
  bool owns_tables = true;
  std::vector<unsigned> grid_host; /* owner only: to build the brick lists of derived plans */
  /* work items per column-segment length: the best length differs by stencil (plan.cu) */
  struct ItemList {
    MarchItem *dev = nullptr;
    int n = 0;
  };
  std::map<unsigned, ItemList> items;
  std::vector<unsigned> boxes; /* kept to build further item lists */
  unsigned *blist_dev = nullptr; /* flat list of every brick in the boxes (naive kernel) */
  long nlist = 0;
  std::map<FieldKey, FieldDev *> fields; /* device arrays, one FieldDev per field */
};

namespace bricks {
/* a table of region links in device memory: links[n] followed by the first 64 KiB chunk of each */
struct LinkBlob {
  bricks_region_link *dev = nullptr;
  int n = 0;
  unsigned nchunks = 0;
};
int link_blob_upload(bricks_ctx *ctx, const bricks_region_link *host, int n, LinkBlob *out);
/* from / to of every link are offsets from src_base / dst_base when that base is not null */
int link_blob_run(bricks_ctx *ctx, const LinkBlob &b, char *src_base, char *dst_base, cudaStream_t s);
void link_blob_free(LinkBlob *b);
}  // namespace bricks

namespace bricks {
/* Peer-to-peer state of a step's exchanges (exchange.cu, "copy-engine exchange"): the peers' field
 * storages and flag words mapped into this process (CUDA IPC between processes, direct pointers
 * between threads of one process), and the sequence number that orders the exchanges of all ranks. */
constexpr int kP2PMaxRanks = 64;
struct P2PState {
  double *field = nullptr;        /* the local storage the mappings belong to */
  unsigned *flags = nullptr;      /* local: ready[kP2PMaxRanks] then done[kP2PMaxRanks], written by the peers */
  char *peer_field[kP2PMaxRanks] = {};
  unsigned *peer_flags[kP2PMaxRanks] = {};
  bool mapped_ipc[kP2PMaxRanks] = {}; /* opened with cudaIpcOpenMemHandle (to close) */
  void *ipc_field[kP2PMaxRanks] = {}, *ipc_flags[kP2PMaxRanks] = {};
  unsigned seq = 0;
  int rank = 0, nranks = 1;
  bool connected = false;
  /* regions of at least ce_min bytes travel as one copy-engine copy each (no SM, ~10 us per copy
   * whatever its size), smaller ones together in ONE store kernel of small CTAs into peer memory:
   * 0 = every region by copy engine, SIZE_MAX = every region in the kernel */
  size_t ce_min = 1u << 20;
};
}  // namespace bricks

struct bricks_exchange {
  bricks_ctx *ctx = nullptr;
  bricks::P2PState *p2p = nullptr; /* borrowed from the step that owns this exchange, or null */
  int rank = 0;
  struct Msg {
    size_t recv_off, send_off, recv_bytes, send_bytes;
    int recv_peer, send_peer;
  };
  std::vector<Msg> msgs;
  bricks::LinkBlob self; /* self neighbours: offsets relative to the storage base */
  bricks::LinkBlob push; /* P2P: this rank's skin regions -> the peers' ghost bricks, absolute addresses */
};

namespace bricks {
/* longest column (brick layers) one warp marches for this stencil; measured on B200
 * (profiles/r01_column_length.txt) */
unsigned column_length(int stencil_id);
/* the plan's item list for that column length (built and uploaded on first use) */
int plan_items(bricks_plan *plan, unsigned maxcol, const MarchItem **items_dev, int *nitems);
int encode_map(CUtensorMap *m, const void *base, unsigned nbricks, size_t step_elems, unsigned bi,
               unsigned bj, unsigned bk, bool swizzle64, unsigned bb = 1, unsigned elem_bytes = 8);
}
