// The sweep schedule of one rank: the hot loop `brick_func` of the reference's MPI drivers
// (weak/main.cu:246-287, strong/main.cu:397-510; CPU twins weak/main.cpp:172-213,
// strong/main.cpp:349-400) as ONE library object, used by bench.py, the tests and the C++ drop-in
// drivers alike:
//
//   comm    : wait(compute) -- [pack] -- exchange -- [unpack] -- [local links] ------------+
//   compute : sweep 1 over the bricks that read no ghost brick (the inner box) ------------+ wait(comm)
//             -- sweep 1 over the shell -- sweeps 2 .. ST_ITER-1 over every brick -- last sweep over
//             the owned bricks only (ghost output is overwritten by the next exchange: the CPU
//             driver's skip, weak/main.cpp:204-209)
//
// EARLY: the last sweep computes the owned skin bricks first, the communication stream starts the
// NEXT step's exchange as soon as they are written, and the rest of that sweep plus the next
// step's first sweep run while the messages are in flight.  The communication program is a list
// of region-copy tables and exchange plans so that the weak driver (exchange straight out of the
// brick storage, or one packed message per peer), the SHIFT exchange (three stages) and the strong
// driver (pack -> per-peer messages -> unpack -> intra-GPU links) share the code.
#include <unistd.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

struct bricks_step {
  bricks_ctx *ctx = nullptr;
  /* box plans over one shared device copy of grid / adj: [0] every brick, [1] the inner box (reads
   * no ghost brick), [2] the shell around it, [3] the owned bricks, [4] the owned skin = owned minus inner */
  bricks_plan *plans[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  bool has_inner = false;
  struct Op {
    int kind; /* 0 copy table, 1 exchange */
    bricks::LinkBlob links;
    int src_base = 0, dst_base = 0;
    bricks_exchange *ex = nullptr;
    int base = 0;
    bool owned = false;
  };
  std::vector<Op> comm;
  char *arena = nullptr;
  size_t arena_bytes = 0;
  const double *in_flight = nullptr; /* field 0 of the storages whose exchange the last step started early */
  /* optional timeline of one step (bricks_step_trace): an event pair around every operation */
  struct Rec {
    int code; /* BRICKS_TRACE_* + detail */
    cudaEvent_t e0, e1;
  };
  bricks::P2PState *p2p = nullptr; /* copy-engine exchange (bricks_step_p2p_*), owned */
  bool trace = false;
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool; /* events are reused between traced steps */
  size_t pool_used = 0;
};

namespace {

/* what a rank publishes about the storage it exchanges (bricks_step_p2p_export): 256 bytes */
struct P2PDesc {
  int pid, device, rank, reserved;
  unsigned long long signature; /* of the region tables: every rank must hold the same ones */
  void *field, *flags;
  cudaIpcMemHandle_t hfield, hflags;
  char pad[256 - 4 * 4 - 8 - 2 * 8 - 2 * sizeof(cudaIpcMemHandle_t)];
};
static_assert(sizeof(P2PDesc) == 256, "P2P descriptor is 256 bytes");

/* the exchanges of the program that move bytes of the field itself with equal send / receive sizes */
bool p2p_eligible(const bricks_step::Op &op) {
  if (op.kind != 1 || op.base != BRICKS_BASE_FIELD) return false;
  for (auto &m : op.ex->msgs)
    if (m.recv_bytes != m.send_bytes) return false;
  return true;
}

unsigned long long p2p_signature(const bricks_step *st) {
  unsigned long long h = 0xCBF29CE484222325ull;
  auto mix = [&](unsigned long long v) {
    h ^= v;
    h *= 0x100000001B3ull;
  };
  for (auto &op : st->comm)
    if (p2p_eligible(op))
      for (auto &m : op.ex->msgs) mix(m.recv_off), mix(m.send_off), mix(m.recv_bytes);
  return h;
}

void p2p_release(bricks_step *st) {
  bricks::P2PState *p = st->p2p;
  if (!p) return;
  for (auto &op : st->comm)
    if (op.kind == 1) {
      op.ex->p2p = nullptr;
      bricks::link_blob_free(&op.ex->push);
    }
  for (int r = 0; r < bricks::kP2PMaxRanks; ++r)
    if (p->mapped_ipc[r]) {
      cudaIpcCloseMemHandle(p->ipc_field[r]);
      cudaIpcCloseMemHandle(p->ipc_flags[r]);
    }
  cudaFree(p->flags);
  delete p;
  st->p2p = nullptr;
}

char *resolve_base(const bricks_step *st, int kind, double *field0) {
  switch (kind) {
    case BRICKS_BASE_FIELD: return (char *)field0;
    case BRICKS_BASE_ARENA: return st->arena;
    default: return nullptr;
  }
}

cudaEvent_t trace_event(bricks_step *st) {
  if (st->pool_used == st->pool.size()) {
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    st->pool.push_back(e);
  }
  return st->pool[st->pool_used++];
}

/* event pair around one operation on `stream` when tracing; fn enqueues the operation */
template <typename F>
int traced(bricks_step *st, int code, int stream, F fn) {
  if (!st->trace) return fn();
  bricks_step::Rec r;
  r.code = code;
  r.e0 = trace_event(st);
  r.e1 = trace_event(st);
  BRK_CUDA(cudaEventRecord(r.e0, st->ctx->streams[stream]));
  if (int rc = fn()) return rc;
  BRK_CUDA(cudaEventRecord(r.e1, st->ctx->streams[stream]));
  st->recs.push_back(r);
  return 0;
}

int run_comm(bricks_step *st, double *field0) {
  bricks_ctx *ctx = st->ctx;
  cudaStream_t s = ctx->streams[BRICKS_STREAM_COMM];
  int idx = 0;
  for (auto &op : st->comm) {
    int rc = 0;
    if (op.kind == 0) {
      rc = traced(st, BRICKS_TRACE_COPY + idx, BRICKS_STREAM_COMM, [&] {
        return bricks::link_blob_run(ctx, op.links, resolve_base(st, op.src_base, field0),
                                     resolve_base(st, op.dst_base, field0), s);
      });
    } else {
      double *base = (double *)resolve_base(st, op.base, field0);
      rc = traced(st, BRICKS_TRACE_EXCHANGE + idx, BRICKS_STREAM_COMM,
                  [&] { return bricks_exchange_run(op.ex, base, BRICKS_STREAM_COMM); });
    }
    if (rc) return rc;
    ++idx;
  }
  return 0;
}

void box6(unsigned *o, unsigned a0, unsigned a1, unsigned a2, unsigned b0, unsigned b1, unsigned b2) {
  o[0] = a0, o[1] = a1, o[2] = a2, o[3] = b0, o[4] = b1, o[5] = b2;
}

/* six disjoint slabs covering [olo, ohi) minus [ilo, ihi), i fastest */
void shell6(unsigned *o, const unsigned olo[3], const unsigned ohi[3], const unsigned ilo[3], const unsigned ihi[3]) {
  box6(o + 0, olo[0], olo[1], olo[2], ohi[0], ohi[1], ilo[2]);
  box6(o + 6, olo[0], olo[1], ihi[2], ohi[0], ohi[1], ohi[2]);
  box6(o + 12, olo[0], olo[1], ilo[2], ohi[0], ilo[1], ihi[2]);
  box6(o + 18, olo[0], ihi[1], ilo[2], ohi[0], ohi[1], ihi[2]);
  box6(o + 24, olo[0], ilo[1], ilo[2], ilo[0], ihi[1], ihi[2]);
  box6(o + 30, ihi[0], ilo[1], ilo[2], ohi[0], ihi[1], ihi[2]);
}

}  // namespace

extern "C" {

int bricks_step_boxes(const unsigned gdims[3], unsigned ghost_bricks, int which, unsigned *boxes, int *nboxes) {
  BRK_ARG(gdims && boxes && nboxes && ghost_bricks >= 1 && which >= 0 && which < 5);
  const unsigned g = ghost_bricks;
  unsigned zero[3] = {0, 0, 0}, full[3], ilo[3], ihi[3], olo[3], ohi[3];
  bool has_inner = true;
  for (int d = 0; d < 3; ++d) {
    BRK_ARG(gdims[d] > 2 * g);
    full[d] = gdims[d];
    olo[d] = g, ohi[d] = gdims[d] - g;
    ilo[d] = 2 * g, ihi[d] = gdims[d] > 4 * g ? gdims[d] - 2 * g : 2 * g;
    if (ihi[d] <= ilo[d]) has_inner = false;
  }
  *nboxes = 0;
  switch (which) {
    case BRICKS_PLAN_FULL: box6(boxes, 0, 0, 0, full[0], full[1], full[2]); *nboxes = 1; break;
    case BRICKS_PLAN_OWNED: box6(boxes, olo[0], olo[1], olo[2], ohi[0], ohi[1], ohi[2]); *nboxes = 1; break;
    case BRICKS_PLAN_INNER:
      if (has_inner) { box6(boxes, ilo[0], ilo[1], ilo[2], ihi[0], ihi[1], ihi[2]); *nboxes = 1; }
      break;
    case BRICKS_PLAN_SHELL:
      if (has_inner) { shell6(boxes, zero, full, ilo, ihi); *nboxes = 6; }
      break;
    case BRICKS_PLAN_OSKIN:
      if (has_inner) { shell6(boxes, olo, ohi, ilo, ihi); *nboxes = 6; }
      break;
  }
  return 0;
}

int bricks_step_create(bricks_ctx *ctx, const unsigned *grid_host, const unsigned gdims[3],
                       const unsigned *adj_host, unsigned nbricks, unsigned ghost_bricks, bricks_step **out) {
  BRK_ARG(ctx && grid_host && gdims && adj_host && out && ghost_bricks >= 1);
  for (int d = 0; d < 3; ++d) BRK_ARG(gdims[d] > 2 * ghost_bricks);
  bricks_step *st = new bricks_step;
  st->ctx = ctx;
  unsigned boxes[36];
  int nb = 0;
  int rc = bricks_step_boxes(gdims, ghost_bricks, BRICKS_PLAN_FULL, boxes, &nb);
  if (!rc) rc = bricks_plan_create(ctx, grid_host, gdims, adj_host, nbricks, nb, boxes, &st->plans[0]);
  for (int which = 1; which < 5 && !rc; ++which) {
    rc = bricks_step_boxes(gdims, ghost_bricks, which, boxes, &nb);
    if (!rc && nb) rc = bricks_plan_derive(st->plans[0], nb, boxes, &st->plans[which]);
  }
  st->has_inner = st->plans[BRICKS_PLAN_INNER] != nullptr;
  if (rc) {
    bricks_step_destroy(st);
    return rc;
  }
  *out = st;
  return 0;
}

int bricks_step_destroy(bricks_step *st) {
  if (!st) return 0;
  cudaSetDevice(st->ctx->device);
  for (int i = 4; i >= 0; --i) bricks_plan_destroy(st->plans[i]); /* derived plans before their parent */
  p2p_release(st);
  for (cudaEvent_t e : st->pool) cudaEventDestroy(e);
  for (auto &op : st->comm) {
    if (op.kind == 0) bricks::link_blob_free(&op.links);
    if (op.kind == 1 && op.owned) bricks_exchange_destroy(op.ex);
  }
  if (st->arena) cudaFree(st->arena);
  delete st;
  return 0;
}

int bricks_step_plan(bricks_step *st, int which, bricks_plan **plan) {
  BRK_ARG(st && plan && which >= 0 && which < 5);
  BRK_ARG(st->plans[which]);
  *plan = st->plans[which];
  return 0;
}

int bricks_step_arena(bricks_step *st, size_t bytes, void **arena_dev) {
  BRK_ARG(st && !st->arena && arena_dev);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  BRK_CUDA(cudaMalloc(&st->arena, bytes ? bytes : 256));
  st->arena_bytes = bytes;
  *arena_dev = st->arena;
  return 0;
}

int bricks_step_add_copy(bricks_step *st, const bricks_region_link *links_host, int n, int src_base, int dst_base) {
  BRK_ARG(st && src_base >= 0 && src_base <= BRICKS_BASE_ARENA && dst_base >= 0 && dst_base <= BRICKS_BASE_ARENA);
  BRK_ARG((src_base != BRICKS_BASE_ARENA && dst_base != BRICKS_BASE_ARENA) || st->arena);
  if (n == 0) return 0;
  BRK_ARG(links_host && n > 0);
  for (int i = 0; i < n; ++i)
    BRK_ARG((links_host[i].len & 15) == 0 && ((uintptr_t)links_host[i].from & 15) == 0 &&
            ((uintptr_t)links_host[i].to & 15) == 0);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  bricks_step::Op op;
  op.kind = 0;
  op.src_base = src_base;
  op.dst_base = dst_base;
  if (int rc = bricks::link_blob_upload(st->ctx, links_host, n, &op.links)) return rc;
  st->comm.push_back(op);
  return 0;
}

int bricks_step_add_exchange(bricks_step *st, bricks_exchange *ex, int base, int take_ownership) {
  BRK_ARG(st && ex && (base == BRICKS_BASE_FIELD || (base == BRICKS_BASE_ARENA && st->arena)));
  bricks_step::Op op;
  op.kind = 1;
  op.ex = ex;
  op.base = base;
  op.owned = take_ownership != 0;
  st->comm.push_back(op);
  return 0;
}

int bricks_step_exchange(bricks_step *st, double *field0) {
  BRK_ARG(st && field0);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  st->in_flight = nullptr;
  if (int rc = bricks_stream_wait(st->ctx, BRICKS_STREAM_COMM, BRICKS_STREAM_COMPUTE)) return rc;
  return run_comm(st, field0);
}

int bricks_step_sweep(bricks_step *st, int which, int stencil_id, int nfields, const double *const *in,
                      double *const *out, const double *coeff_host, int stream) {
  BRK_ARG(st && which >= 0 && which < 5 && st->plans[which]);
  return bricks_stencil(st->plans[which], stencil_id, nfields, in, out, 512, 512, coeff_host, stream);
}

int bricks_step_p2p_export(bricks_step *st, double *field0, int rank, void *desc256) {
  BRK_ARG(st && field0 && desc256 && rank >= 0 && rank < bricks::kP2PMaxRanks);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  p2p_release(st);
  bricks::P2PState *p = new bricks::P2PState;
  st->p2p = p;
  p->field = field0;
  p->rank = rank;
  const size_t fb = 2 * bricks::kP2PMaxRanks * sizeof(unsigned);
  BRK_CUDA(cudaMalloc(&p->flags, fb));
  BRK_CUDA(cudaMemset(p->flags, 0, fb));
  BRK_CUDA(cudaDeviceSynchronize());
  P2PDesc d;
  memset(&d, 0, sizeof d);
  d.pid = (int)getpid();
  d.device = st->ctx->device;
  d.rank = rank;
  d.signature = p2p_signature(st);
  d.field = field0;
  d.flags = p->flags;
  /* handles of the whole allocations: field0 must be the start of one (cudaMalloc / bricks_malloc) */
  BRK_CUDA(cudaIpcGetMemHandle(&d.hfield, field0));
  BRK_CUDA(cudaIpcGetMemHandle(&d.hflags, p->flags));
  memcpy(desc256, &d, sizeof d);
  return 0;
}

int bricks_step_p2p_disable(bricks_step *st) {
  BRK_ARG(st);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  p2p_release(st);
  return 0;
}

int bricks_step_p2p_connect(bricks_step *st, int nranks, const void *descs) {
  BRK_ARG(st && st->p2p && descs && nranks >= 1 && nranks <= bricks::kP2PMaxRanks);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  {
    /* the exchange is ordered by stream memory operations: refuse here, where the caller can still
     * fall back to NCCL on every rank, not at the first step */
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    for (const char *name : {"cuStreamWaitValue32", "cuStreamWriteValue32"})
      if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        bricks::set_error("bricks_step_p2p_connect: %s is not available from this driver", name);
        return -6;
      }
  }
  bricks::P2PState *p = st->p2p;
  p->nranks = nranks;
  const P2PDesc *d = (const P2PDesc *)descs;
  const unsigned long long sig = p2p_signature(st);
  bool peer[bricks::kP2PMaxRanks] = {};
  int eligible = 0;
  for (auto &op : st->comm) {
    if (!p2p_eligible(op)) continue;
    ++eligible;
    for (auto &m : op.ex->msgs) {
      BRK_ARG(m.recv_peer >= 0 && m.recv_peer < nranks && m.send_peer >= 0 && m.send_peer < nranks);
      if (m.recv_peer != p->rank) peer[m.recv_peer] = true;
      if (m.send_peer != p->rank) peer[m.send_peer] = true;
    }
  }
  if (!eligible) {
    bricks::set_error("bricks_step_p2p_connect: the communication program has no exchange of the field itself "
                      "(packed exchanges go through an arena and stay on NCCL)");
    return -1;
  }
  for (int r = 0; r < nranks; ++r) {
    if (!peer[r]) continue;
    if (d[r].rank != r || d[r].signature != sig) {
      bricks::set_error("bricks_step_p2p_connect: rank %d holds other region tables (unequal subdomains?)", r);
      return -1;
    }
    if (d[r].pid == (int)getpid()) { /* a thread of this process: its pointers are valid here */
      if (d[r].device != st->ctx->device) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(d[r].device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else BRK_CUDA(e);
      }
      p->peer_field[r] = (char *)d[r].field;
      p->peer_flags[r] = (unsigned *)d[r].flags;
    } else {
      BRK_CUDA(cudaIpcOpenMemHandle(&p->ipc_field[r], d[r].hfield, cudaIpcMemLazyEnablePeerAccess));
      BRK_CUDA(cudaIpcOpenMemHandle(&p->ipc_flags[r], d[r].hflags, cudaIpcMemLazyEnablePeerAccess));
      p->mapped_ipc[r] = true;
      p->peer_field[r] = (char *)p->ipc_field[r];
      p->peer_flags[r] = (unsigned *)p->ipc_flags[r];
    }
  }
  /* experiments: BRICKS_P2P_MODE=ce (every region by copy engine) / push (every region in the store kernel) */
  const char *mode = getenv("BRICKS_P2P_MODE");
  if (mode && strcmp(mode, "ce") == 0) p->ce_min = 0;
  if (mode && strcmp(mode, "push") == 0) p->ce_min = (size_t)-1;
  if (const char *e = getenv("BRICKS_P2P_CE_MIN")) /* bytes; tests lower it so small subdomains take both paths */
    if (atoll(e) > 0) p->ce_min = (size_t)atoll(e);
  for (auto &op : st->comm) {
    if (!p2p_eligible(op)) continue;
    std::vector<bricks_region_link> links;
    for (auto &m : op.ex->msgs)
      if (m.send_peer != p->rank && m.send_bytes && m.send_bytes < p->ce_min) {
        bricks_region_link l;
        l.from = (double *)((char *)p->field + m.send_off);
        l.to = (double *)(p->peer_field[m.send_peer] + m.recv_off);
        l.len = m.send_bytes;
        links.push_back(l);
      }
    if (!links.empty())
      if (int rc = bricks::link_blob_upload(st->ctx, links.data(), (int)links.size(), &op.ex->push)) return rc;
    op.ex->p2p = p;
  }
  p->connected = true;
  return 0;
}

int bricks_step_trace(bricks_step *st, int enable) {
  BRK_ARG(st);
  st->trace = enable != 0;
  return 0;
}

int bricks_step_trace_read(bricks_step *st, int cap, int *n, int *codes, float *start_ms, float *end_ms) {
  BRK_ARG(st && n);
  *n = (int)st->recs.size();
  if (cap <= 0 || st->recs.empty()) return 0;
  BRK_ARG(codes && start_ms && end_ms && cap >= *n);
  BRK_CUDA(cudaSetDevice(st->ctx->device));
  BRK_CUDA(cudaDeviceSynchronize());
  /* origin: the earliest start among the records */
  cudaEvent_t origin = st->recs[0].e0;
  for (auto &r : st->recs) {
    float d = 0;
    BRK_CUDA(cudaEventElapsedTime(&d, origin, r.e0));
    if (d < 0) origin = r.e0;
  }
  for (size_t i = 0; i < st->recs.size(); ++i) {
    codes[i] = st->recs[i].code;
    BRK_CUDA(cudaEventElapsedTime(&start_ms[i], origin, st->recs[i].e0));
    BRK_CUDA(cudaEventElapsedTime(&end_ms[i], origin, st->recs[i].e1));
  }
  return 0;
}

int bricks_step_reset(bricks_step *st) {
  BRK_ARG(st);
  st->in_flight = nullptr;
  return 0;
}

int bricks_step_run(bricks_step *st, int stencil_id, int st_iter, int nfields, double *const *a,
                    double *const *b, const double *coeff_host, int flags) {
  BRK_ARG(st && a && b && nfields > 0 && st_iter >= 2 && st_iter % 2 == 0);
  bricks_ctx *ctx = st->ctx;
  const int C = BRICKS_STREAM_COMPUTE, X = BRICKS_STREAM_COMM;
  BRK_CUDA(cudaSetDevice(ctx->device));
  if (st->trace) { /* the timeline holds the last traced step */
    st->recs.clear();
    st->pool_used = 0;
  }
  const bool overlap = (flags & BRICKS_STEP_OVERLAP) && st->has_inner;
  const bool early = overlap && (flags & BRICKS_STEP_EARLY);
  int rc = 0;
  /* the exchange reads the skin written by the previous step's last sweep -- unless that step
   * already started it behind its own last sweep (EARLY) */
  const bool arrived_early = st->in_flight == a[0];
  const bool split_first = overlap && !(arrived_early && (flags & BRICKS_STEP_EARLY_UNSPLIT));
  /* The communication program is ordered behind the compute stream HERE (event recorded now), but
   * where a sweep runs beside it that sweep is enqueued first: an exchange can be ~50 host calls
   * (one copy per region), and the sweep it hides behind should not wait for them to be issued. */
  bool comm_pending = false;
  if (!arrived_early) {
    if ((rc = bricks_stream_wait(ctx, X, C))) return rc;
    if (split_first)
      comm_pending = true;
    else if ((rc = run_comm(st, a[0])))
      return rc;
  }
  st->in_flight = nullptr;
  double *const *src = a, *const *dst = b;
  for (int it = 0; it < st_iter; ++it) {
    const bool last = it == st_iter - 1;
    auto sweep = [&](int which) {
      return traced(st, BRICKS_TRACE_SWEEP + 10 * it + which, C, [&] {
        return bricks_stencil(st->plans[which], stencil_id, nfields, src, dst, 512, 512, coeff_host, C);
      });
    };
    if (it == 0) {
      if (split_first) {
        if ((rc = sweep(1))) return rc;
        if (comm_pending && (rc = run_comm(st, a[0]))) return rc;
        if ((rc = bricks_stream_wait(ctx, C, X))) return rc;
        if ((rc = sweep(2))) return rc;
      } else {
        /* serial schedule, or the messages have had the whole inner part of the previous step's
         * last sweep to arrive: one unsplit sweep instead of inner + shell */
        if ((rc = bricks_stream_wait(ctx, C, X))) return rc;
        if ((rc = sweep(0))) return rc;
      }
    } else if (last && early) {
      if ((rc = sweep(4))) return rc;
      if ((rc = bricks_stream_wait(ctx, X, C))) return rc; /* behind the skin bricks, not behind the inner box */
      if ((rc = sweep(1))) return rc;
      if ((rc = run_comm(st, dst[0]))) return rc; /* dst == a: the input of the next step */
      st->in_flight = dst[0];
    } else if (last) {
      if ((rc = sweep(3))) return rc;
    } else {
      if ((rc = sweep(0))) return rc;
    }
    double *const *t = src;
    src = dst;
    dst = t;
  }
  return 0;
}

}  // extern "C"
