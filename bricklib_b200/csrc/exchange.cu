// Region copies and the halo exchange.
//   bricks_copy_regions  replaces cudaCopy<<<nlinks,64>>> (strong/main.cu:62-83): the reference
//                        gives each region ONE block of 64 threads moving 8-byte words; here a
//                        region is split over as many CTAs as its size warrants, 16 bytes/lane.
//   bricks_exchange_*    replaces BrickDecomp::exchange (include/brick-mpi.h:466-495) and the
//                        D2H -> MPI -> H2D staging of weak/main.cu:251-275 with device-resident
//                        ncclSend/ncclRecv straight from/to brick storage (regions are contiguous
//                        by construction of the layout, so nothing is packed); regions whose
//                        peer is this rank become on-device copies.
#include <alloca.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace {

constexpr int kCopyThreads = 256;
constexpr size_t kChunk = 64 * 1024; /* bytes per CTA */

/* links[] in device memory; grid.x = total chunks.  A non-null base turns the `from` (src_base) or
 * `to` (dst_base) members into byte offsets from it: exchange plans are built once and run on
 * whatever storage the step is given (self links: both ends in the field; pack / unpack: one end
 * in the field, the other in the message arena). */
__global__ void __launch_bounds__(kCopyThreads) copy_regions_kernel(const bricks_region_link *links,
                                                                   const unsigned *chunk_first, int nlinks,
                                                                   char *src_base, char *dst_base) {
  /* binary search: which link does this CTA's chunk belong to */
  int lo = 0, hi = nlinks - 1;
  const unsigned c = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (chunk_first[mid] <= c) lo = mid; else hi = mid - 1;
  }
  const bricks_region_link l = links[lo];
  const size_t start = (size_t)(c - chunk_first[lo]) * kChunk;
  const size_t end = start + kChunk < l.len ? start + kChunk : l.len;
  const char *src = (src_base ? src_base + (size_t)l.from : (const char *)l.from) + start;
  char *dst = (dst_base ? dst_base + (size_t)l.to : (char *)l.to) + start;
  const size_t n16 = (end - start) >> 4;
  const int4 *s4 = reinterpret_cast<const int4 *>(src);
  int4 *d4 = reinterpret_cast<int4 *>(dst);
  for (size_t i = threadIdx.x; i < n16; i += 4 * kCopyThreads) {
    int4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * kCopyThreads < n16) v[u] = __ldcs(s4 + i + u * kCopyThreads);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * kCopyThreads < n16) __stcs(d4 + i + u * kCopyThreads, v[u]);
  }
}

/* Scratch of bricks_copy_regions_static: the link table + first-chunk table of ONE call, in device
 * memory that exists from library load to process exit on every GPU that loads the library
 * (kStaticMaxLinks * 28 bytes = 56 KiB).  In host code the name below is only a stand-in for the
 * symbol: the device address comes from cudaGetSymbolAddress, once per device (static_table()). */
constexpr int kStaticMaxLinks = BRICKS_COPY_STATIC_MAX_LINKS;
struct StaticLinkTable {
  bricks_region_link links[kStaticMaxLinks];
  unsigned chunk_first[kStaticMaxLinks];
};
static __device__ __align__(256) StaticLinkTable g_static_links;

/* device address of g_static_links on the CURRENT device, looked up once per device */
int static_table(int device, StaticLinkTable **out) {
  static std::mutex m;
  static StaticLinkTable *addr[64] = {nullptr};
  BRK_ARG(device >= 0 && device < 64);
  std::lock_guard<std::mutex> lock(m);
  if (!addr[device]) {
    void *p = nullptr;
    BRK_CUDA(cudaGetSymbolAddress(&p, g_static_links)); /* the caller made `device` current */
    addr[device] = (StaticLinkTable *)p;
  }
  *out = addr[device];
  return 0;
}

}  // namespace

namespace bricks {

int link_blob_upload(bricks_ctx *ctx, const bricks_region_link *host, int n, LinkBlob *out) {
  /* device blob = links[n] followed by chunk_first[n] */
  std::vector<unsigned> first(n);
  unsigned tot = 0;
  for (int i = 0; i < n; ++i) {
    first[i] = tot;
    tot += (unsigned)((host[i].len + kChunk - 1) / kChunk);
  }
  const size_t lb = sizeof(bricks_region_link) * n, fb = sizeof(unsigned) * n;
  char *d = nullptr;
  BRK_CUDA(cudaMalloc(&d, lb + fb + 16));
  cudaStream_t s = ctx->streams[BRICKS_STREAM_COMPUTE];
  BRK_CUDA(cudaMemcpyAsync(d, host, lb, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyAsync(d + lb, first.data(), fb, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaStreamSynchronize(s));
  out->dev = (bricks_region_link *)d;
  out->n = n;
  out->nchunks = tot;
  return 0;
}

int link_blob_run(bricks_ctx *ctx, const LinkBlob &b, char *src_base, char *dst_base, cudaStream_t s) {
  if (!b.dev || b.nchunks == 0) return 0;
  const unsigned *first = (const unsigned *)((const char *)b.dev + sizeof(bricks_region_link) * b.n);
  copy_regions_kernel<<<b.nchunks, kCopyThreads, 0, s>>>(b.dev, first, b.n, src_base, dst_base);
  BRK_CUDA(cudaGetLastError());
  ++g_launches;
  return 0;
}

void link_blob_free(LinkBlob *b) {
  if (b->dev) cudaFree(b->dev);
  *b = LinkBlob();
}

}  // namespace bricks

/* Tables handed out through the C ABI (bricks_links_upload) are plain device pointers; their
 * sizes are kept host side, keyed by the pointer. */
static std::map<const void *, bricks::LinkBlob> g_link_meta;
static std::mutex g_link_mutex; /* one host thread per device may be inside the library at once */

static bool meta_get(const void *blob, bricks::LinkBlob *out) {
  std::lock_guard<std::mutex> lock(g_link_mutex);
  auto it = g_link_meta.find(blob);
  if (it == g_link_meta.end()) return false;
  *out = it->second;
  return true;
}

extern "C" {

int bricks_links_upload(bricks_ctx *ctx, const bricks_region_link *links_host, int n,
                        bricks_region_link **links_dev) {
  BRK_ARG(ctx && links_host && links_dev && n > 0);
  for (int i = 0; i < n; ++i)
    BRK_ARG((links_host[i].len & 15) == 0 && ((uintptr_t)links_host[i].from & 15) == 0 &&
            ((uintptr_t)links_host[i].to & 15) == 0);
  BRK_CUDA(cudaSetDevice(ctx->device));
  bricks::LinkBlob b;
  if (int rc = bricks::link_blob_upload(ctx, links_host, n, &b)) return rc;
  {
    std::lock_guard<std::mutex> lock(g_link_mutex);
    g_link_meta[b.dev] = b;
  }
  *links_dev = b.dev;
  return 0;
}

int bricks_links_free(bricks_ctx *ctx, bricks_region_link *links_dev) {
  BRK_ARG(ctx);
  if (!links_dev) return 0;
  bricks::LinkBlob b;
  BRK_ARG(meta_get(links_dev, &b)); /* only tables made by bricks_links_upload */
  {
    std::lock_guard<std::mutex> lock(g_link_mutex);
    g_link_meta.erase(links_dev);
  }
  BRK_CUDA(cudaSetDevice(ctx->device));
  BRK_CUDA(cudaFree(links_dev));
  return 0;
}

int bricks_copy_regions(bricks_ctx *ctx, const bricks_region_link *links_dev, int n, int stream) {
  BRK_ARG(ctx && links_dev && stream >= 0 && stream < bricks::kNumStreams);
  bricks::LinkBlob b;
  BRK_ARG(meta_get(links_dev, &b) && b.n == n);
  BRK_CUDA(cudaSetDevice(ctx->device));
  return bricks::link_blob_run(ctx, b, nullptr, nullptr, ctx->streams[stream]);
}

int bricks_copy_regions_static(bricks_ctx *ctx, const bricks_region_link *links_host, int n, int stream) {
  BRK_ARG(ctx && links_host && stream >= 0 && stream < bricks::kNumStreams);
  if (n < 0 || n > kStaticMaxLinks) {
    bricks::set_error("bricks_copy_regions_static: %d links, the static table holds at most %d "
                      "(BRICKS_COPY_STATIC_MAX_LINKS); use bricks_links_upload + bricks_copy_regions", n, kStaticMaxLinks);
    return -1;
  }
  if (n == 0) return 0;
  /* the table of this call, built on the stack: no host allocation either */
  StaticLinkTable *host = (StaticLinkTable *)alloca(sizeof(bricks_region_link) * n + sizeof(unsigned) * n);
  bricks_region_link *hl = (bricks_region_link *)host;
  unsigned *hf = (unsigned *)(hl + n);
  unsigned long long tot = 0;
  for (int i = 0; i < n; ++i) {
    BRK_ARG((links_host[i].len & 15) == 0 && ((uintptr_t)links_host[i].from & 15) == 0 &&
            ((uintptr_t)links_host[i].to & 15) == 0);
    hl[i] = links_host[i];
    hf[i] = (unsigned)tot;
    tot += (links_host[i].len + kChunk - 1) / kChunk;
  }
  BRK_ARG(tot <= 0x7fffffffull);
  if (tot == 0) return 0;
  BRK_CUDA(cudaSetDevice(ctx->device));
  StaticLinkTable *tab = nullptr;
  if (int rc = static_table(ctx->device, &tab)) return rc;
  cudaStream_t s = ctx->streams[stream];
  /* stream-ordered behind earlier work of THIS stream (an earlier call's kernel included); the
   * entries past n keep the last call's contents and are never read.  Pageable source: the call
   * returns once the bytes are staged, so the stack table may die at return. */
  BRK_CUDA(cudaMemcpyAsync(tab->links, hl, sizeof(bricks_region_link) * n, cudaMemcpyHostToDevice, s));
  BRK_CUDA(cudaMemcpyAsync(tab->chunk_first, hf, sizeof(unsigned) * n, cudaMemcpyHostToDevice, s));
  copy_regions_kernel<<<(unsigned)tot, kCopyThreads, 0, s>>>(tab->links, tab->chunk_first, n, nullptr, nullptr);
  BRK_CUDA(cudaGetLastError());
  ++bricks::g_launches;
  return 0;
}

int bricks_comm_unique_id(void *id128) {
  BRK_ARG(id128);
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  BRK_NCCL(ncclGetUniqueId((ncclUniqueId *)id128));
  return 0;
}

int bricks_comm_init_config(bricks_ctx *ctx, int rank, int nranks, const void *id128, int max_ctas) {
  BRK_ARG(ctx && id128 && nranks > 0 && rank >= 0 && rank < nranks && max_ctas >= 0);
  BRK_CUDA(cudaSetDevice(ctx->device));
  ncclUniqueId id;
  memcpy(&id, id128, sizeof id);
  if (max_ctas > 0) {
    /* the exchange runs beside a sweep kernel that fills every SM with one-warp CTAs; a capped
     * send/recv kernel takes fewer SM slots away from it (NVLink needs few CTAs for 100 MB) */
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    cfg.minCTAs = 1;
    cfg.maxCTAs = max_ctas;
    BRK_NCCL(ncclCommInitRankConfig(&ctx->comm, nranks, id, rank, &cfg));
  } else {
    BRK_NCCL(ncclCommInitRank(&ctx->comm, nranks, id, rank));
  }
  ctx->rank = rank;
  ctx->nranks = nranks;
  return 0;
}

int bricks_comm_init(bricks_ctx *ctx, int rank, int nranks, const void *id128) {
  return bricks_comm_init_config(ctx, rank, nranks, id128, 0);
}

int bricks_comm_destroy(bricks_ctx *ctx) {
  BRK_ARG(ctx);
  if (ctx->comm) {
    BRK_NCCL(ncclCommDestroy(ctx->comm));
    ctx->comm = nullptr;
  }
  return 0;
}

int bricks_exchange_create_v(bricks_ctx *ctx, int rank, int nregions, const size_t *recv_off,
                             const size_t *recv_bytes, const int *recv_peer, const size_t *send_off,
                             const size_t *send_bytes, const int *send_peer, bricks_exchange **out) {
  BRK_ARG(ctx && out && nregions > 0 && recv_off && send_off && recv_bytes && send_bytes && recv_peer && send_peer);
  BRK_CUDA(cudaSetDevice(ctx->device));
  std::vector<bricks_exchange::Msg> msgs;
  for (int i = 0; i < nregions; ++i) {
    BRK_ARG((recv_off[i] & 15) == 0 && (send_off[i] & 15) == 0 && (recv_bytes[i] & 15) == 0 &&
            (send_bytes[i] & 15) == 0);
    /* a self neighbour pairs this rank's own send i with its own receive i */
    BRK_ARG((recv_peer[i] == rank) == (send_peer[i] == rank));
    BRK_ARG(recv_peer[i] != rank || recv_bytes[i] == send_bytes[i]);
    if (recv_bytes[i] == 0 && send_bytes[i] == 0) continue;
    bricks_exchange::Msg m = {recv_off[i], send_off[i], recv_bytes[i], send_bytes[i], recv_peer[i], send_peer[i]};
    msgs.push_back(m);
  }
  bricks_exchange *ex = new bricks_exchange;
  ex->ctx = ctx;
  ex->rank = rank;
  ex->msgs = msgs;
  /* Region i is received from recv_peer[i] and sent to send_peer[i].  When recv_peer[i] is this
   * rank, the matching send is this rank's own region i (periodic process grid of extent 1 in
   * that direction), i.e. ghost[i] <- skin[i] on the device. */
  std::vector<bricks_region_link> self;
  for (auto &m : ex->msgs)
    if (m.recv_peer == rank && m.recv_bytes) {
      bricks_region_link l;
      l.from = (double *)m.send_off; /* relative to the storage base */
      l.to = (double *)m.recv_off;
      l.len = m.recv_bytes;
      self.push_back(l);
    }
  if (!self.empty()) {
    int rc = bricks::link_blob_upload(ctx, self.data(), (int)self.size(), &ex->self);
    if (rc) {
      delete ex;
      return rc;
    }
  }
  *out = ex;
  return 0;
}

int bricks_exchange_create(bricks_ctx *ctx, int rank, int nregions, const size_t *recv_off,
                           const size_t *send_off, const size_t *nbytes, const int *recv_peer,
                           const int *send_peer, bricks_exchange **out) {
  return bricks_exchange_create_v(ctx, rank, nregions, recv_off, nbytes, recv_peer, send_off, nbytes, send_peer, out);
}

int bricks_exchange_destroy(bricks_exchange *ex) {
  if (!ex) return 0;
  cudaSetDevice(ex->ctx->device);
  bricks::link_blob_free(&ex->self);
  bricks::link_blob_free(&ex->push);
  delete ex;
  return 0;
}

}  // extern "C"

namespace {

typedef CUresult (*StreamValueFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
StreamValueFn lookup_stream_op(const char *name) {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
    return (StreamValueFn)p;
  return nullptr;
}

/* The peer-memory exchange.  NCCL's send/recv kernel needs most of an SM per CTA and runs beside a
 * sweep that keeps every SM full of one-warp CTAs: measured, an exchange of 104 MB that NVLink moves
 * in ~0.15 ms takes 0.6 ms (2 GPUs) to 0.8 ms (8 GPUs) next to a sweep (profiles/r03_scaling/).
 * Here this rank's skin regions go straight into the peers' ghost bricks (peer-mapped memory over
 * NVLink): each large region as ONE cudaMemcpyAsync on the copy engines, the small ones together in
 * one store kernel, ordered by two flag words per peer pair that carry the exchange's sequence number
 * (stream memory operations):
 *   receiver: ready[me] := seq in every sender's memory   -- my ghost bricks may be overwritten
 *   sender  : wait ready[peer] >= seq for all receivers; move regions; done[me] := seq in each receiver's memory
 *   receiver: wait done[peer] >= seq for every sender
 * Every rank writes all its ready flags before its first wait, so no cycle of waits can form.
 * All ranks hold the same region tables (equal subdomains), so region i lands at the receiver's
 * recv_off[i] = this rank's recv_off[i]. */
int exchange_run_p2p(bricks_exchange *ex, cudaStream_t s) {
  static const StreamValueFn wait_fn = lookup_stream_op("cuStreamWaitValue32");
  static const StreamValueFn write_fn = lookup_stream_op("cuStreamWriteValue32");
  if (!wait_fn || !write_fn) {
    bricks::set_error("stream memory operations are not available from this driver");
    return -6;
  }
  bricks::P2PState *p = ex->p2p;
  const unsigned seq = ++p->seq;
  const int me = ex->rank;
  bool recv_from[bricks::kP2PMaxRanks] = {}, send_to[bricks::kP2PMaxRanks] = {};
  for (auto &m : ex->msgs) {
    if (m.recv_peer != me && m.recv_bytes) recv_from[m.recv_peer] = true;
    if (m.send_peer != me && m.send_bytes) send_to[m.send_peer] = true;
  }
  auto cu = [](CUresult r, const char *what) {
    if (r == CUDA_SUCCESS) return 0;
    bricks::set_error("%s failed with CUresult %d", what, (int)r);
    return 3000 + (int)r;
  };
  /* 1. my ghost regions are free: tell every rank that sends to me */
  for (int r = 0; r < p->nranks; ++r)
    if (recv_from[r])
      if (int rc = cu(write_fn(s, (CUdeviceptr)(p->peer_flags[r] + me), seq, CU_STREAM_WRITE_VALUE_DEFAULT), "cuStreamWriteValue32"))
        return rc;
  /* 2. once every receiver is ready: the large regions by copy engine (no SM involved; a copy costs
   * ~10 us whatever its size, so 42 copies to 7 peers took 0.7 ms, profiles/r03_scaling/p2p_n8), the
   * many small ones in ONE kernel of 256-thread CTAs that stores into the peers' memory (they
   * interleave with the sweep's one-warp CTAs; all bytes that way cost the sweep 12 %,
   * profiles/r03_scaling/p2p_modes_n2) */
  for (int r = 0; r < p->nranks; ++r)
    if (send_to[r])
      if (int rc = cu(wait_fn(s, (CUdeviceptr)(p->flags + r), seq, CU_STREAM_WAIT_VALUE_GEQ), "cuStreamWaitValue32")) return rc;
  char *base = (char *)p->field;
  for (auto &m : ex->msgs)
    if (m.send_peer != me && m.send_bytes && m.send_bytes >= p->ce_min)
      BRK_CUDA(cudaMemcpyAsync(p->peer_field[m.send_peer] + m.recv_off, base + m.send_off, m.send_bytes,
                               cudaMemcpyDeviceToDevice, s));
  if (int rc = bricks::link_blob_run(ex->ctx, ex->push, nullptr, nullptr, s)) return rc;
  for (int r = 0; r < p->nranks; ++r)
    if (send_to[r])
      if (int rc = cu(write_fn(s, (CUdeviceptr)(p->peer_flags[r] + bricks::kP2PMaxRanks + me), seq, CU_STREAM_WRITE_VALUE_DEFAULT),
                      "cuStreamWriteValue32"))
        return rc;
  /* 3. everything bound for me has landed */
  for (int r = 0; r < p->nranks; ++r)
    if (recv_from[r])
      if (int rc = cu(wait_fn(s, (CUdeviceptr)(p->flags + bricks::kP2PMaxRanks + r), seq, CU_STREAM_WAIT_VALUE_GEQ), "cuStreamWaitValue32"))
        return rc;
  return 0;
}

}  // namespace

extern "C" {

/* Stream order is the only ordering between two runs of the same plan (and between a run and the
 * sweeps that read what it delivered): callers enqueue bricks_stream_wait between the streams. */
int bricks_exchange_run(bricks_exchange *ex, double *storage, int stream) {
  BRK_ARG(ex && storage && stream >= 0 && stream < bricks::kNumStreams);
  bricks_ctx *ctx = ex->ctx;
  BRK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->streams[stream];
  char *base = (char *)storage;
  bool remote = false;
  for (auto &m : ex->msgs) remote = remote || m.recv_peer != ex->rank || m.send_peer != ex->rank;
  if (remote && ex->p2p && ex->p2p->connected && ex->p2p->field == storage) {
    /* the peers' copies of THIS storage are mapped here: copy engines over NVLink, no NCCL kernel */
    if (int rc = exchange_run_p2p(ex, s)) return rc;
  } else if (remote) {
    if (!ctx->comm) {
      bricks::set_error("exchange has remote peers but bricks_comm_init was not called");
      return -4;
    }
    /* both sides enumerate regions in the same order i (the reference's MPI tag, brick-mpi.h:478-484);
     * NCCL matches send/recv pairs per peer in issue order */
    BRK_NCCL(ncclGroupStart());
    for (auto &m : ex->msgs) {
      if (m.recv_peer != ex->rank && m.recv_bytes)
        BRK_NCCL(ncclRecv(base + m.recv_off, m.recv_bytes, ncclChar, m.recv_peer, ctx->comm, s));
      if (m.send_peer != ex->rank && m.send_bytes)
        BRK_NCCL(ncclSend(base + m.send_off, m.send_bytes, ncclChar, m.send_peer, ctx->comm, s));
    }
    BRK_NCCL(ncclGroupEnd());
  }
  return bricks::link_blob_run(ctx, ex->self, base, base, s);
}

}  // extern "C"
