// strong.cpp -- strong-scaling driver: a fixed dom^3 periodic grid in (dom/sdom)^3 subdomains named
// by Z-Morton ids, contiguous id ranges per GPU.  Drop-in for the reference's strong/main.cu
// (flow :113-574) and strong/args.cpp (flags -d -s -I -v -h), with -S 7|13|25|125 for the
// stencil, -G ngpus replacing `mpirun -np` and -n for the serial schedule.
//
// Per step (strong/main.cu:397-510): pack the skins bound for other GPUs, one message per peer
// (NCCL), unpack into the ghost bricks, device copies between subdomains of the same GPU, then
// ST_ITER sweeps, each ONE launch batched over all local subdomains.  Prints the reference's
// lines (strong/main.cu:550-566).
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>

#include "bricks/multiarray.h"
#include "bricks/strong-plan.h"
#include "common.h"

namespace {
constexpr unsigned GZ = 8, TILE = 8;
typedef BrickDecomp<3, 8, 8, 8> Decomp3D;
typedef BrickStencilPlan::Box Box;

const char *help =
    "Running the strong-scaling brick stencil on B200s with %s\n\n"
    "Program options\n"
    "  -h: show help (this message)\n"
    "  -G: number of GPUs (ranks), default 1\n"
    "  -d: Int, overall domain size is d^3 (default 512)\n"
    "  -s: Int, subdomain size is s^3 (default 128)\n"
    "  -I: number of iterations, default %d\n"
    "  -S: stencil 7|13|25|125, default 125\n"
    "  -v: validate against the naive device kernel\n"
    "  -n: no overlap (exchange, then sweeps)\n"
    "Example usage:\n"
    "  %s -G 8 -d 1024 -s 128\n";

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

/// a RegionLink list on the device (strong/main.cu:62-83 cudaCopy)
struct Links {
  bricks_ctx *ctx = nullptr;
  bricks_region_link *dev = nullptr;
  int n = 0;
  void upload(bricks_ctx *c, const std::vector<bricks_region_link> &host) {
    ctx = c, n = (int)host.size();
    if (n) gpuCheck(bricks_links_upload(c, host.data(), n, &dev));
  }
  void run(int stream) const {
    if (n) gpuCheck(bricks_copy_regions(ctx, dev, n, stream));
  }
  ~Links() {
    if (dev) bricks_links_free(ctx, dev);
  }
};

std::vector<Box> shell_of(const unsigned g[3], const unsigned lo[3], const unsigned hi[3]) {
  auto box = [](unsigned a0, unsigned a1, unsigned a2, unsigned b0, unsigned b1, unsigned b2) {
    Box b = {{a0, a1, a2}, {b0, b1, b2}};
    return b;
  };
  return {box(0, 0, 0, g[0], g[1], lo[2]),           box(0, 0, hi[2], g[0], g[1], g[2]),
          box(0, 0, lo[2], g[0], lo[1], hi[2]),      box(0, hi[1], lo[2], g[0], g[1], hi[2]),
          box(0, lo[1], lo[2], lo[0], hi[1], hi[2]), box(hi[0], lo[1], lo[2], g[0], hi[1], hi[2])};
}
}  // namespace

int main(int argc, char **argv) {
  unsigned dom_size = 512, sdom_size = 128;
  int iters = 100, ngpus = 1, which = 125, c;
  bool validate = false, overlap = true;
  while ((c = getopt(argc, argv, "d:s:I:vhG:S:n")) != -1) {
    switch (c) {
      case 'd': dom_size = std::stoi(optarg); break;
      case 's': sdom_size = std::stoi(optarg); break;
      case 'I': iters = std::stoi(optarg); break;
      case 'v': validate = true; break;
      case 'G': ngpus = std::stoi(optarg); break;
      case 'S': which = std::stoi(optarg); break;
      case 'n': overlap = false; break;
      default: printf("Unknown options %c\n", c);  // fall through
      case 'h': printf(help, argv[0], iters, argv[0]); return 0;
    }
  }
  if (!drv::stencil_known(which) || ngpus < 1 || iters < 1 || sdom_size < 2 * TILE || sdom_size % TILE) {
    printf("Unknown stencil / bad -G / bad -I / bad -s\n");
    return 1;
  }
  const int st_iter = drv::stencil_iters(which);
  {
    long n = dom_size / sdom_size;
    n = n * n * n;
    std::cout << "Pagesize " << sysconf(_SC_PAGESIZE) << "; MPI Size " << ngpus << " * OpenMP threads 1" << std::endl;
    std::cout << "Domain size of " << dom_size << "^3 decomposed into " << sdom_size << "^3 subdomains" << std::endl;
    std::cout << "Total of " << n << " subdomains, " << (n + ngpus - 1) / ngpus << " per rank" << std::endl;
  }
  int failures = 0;
  std::mutex fail_mutex;

  int rc = drv::run_ranks(ngpus, 10, [&](drv::Rank &rk) {
    // all subdomains share one decomposition / BrickInfo (strong/main.cu:131-153)
    Decomp3D bDecomp({sdom_size, sdom_size, sdom_size}, GZ);
    bDecomp.initialize(skin3d_good);
    BrickInfo<3> bInfo = bDecomp.getBrickInfo();
    const unsigned g[3] = {bDecomp.gridDim(0), bDecomp.gridDim(1), bDecomp.gridDim(2)};
    const unsigned *grid_ptr = bDecomp.gridData();
    for (unsigned k = 1; k + 1 < g[2]; ++k)
      for (unsigned j = 1; j + 1 < g[1]; ++j)
        for (unsigned i = 1; i + 1 < g[0]; ++i) {
          const unsigned l = grid_ptr[(k * g[1] + j) * g[0] + i];
          for (unsigned id = 0; id < 27; ++id)
            if (bInfo.adj[bInfo.adj[l][id]][26 - id] != l) throw std::runtime_error("err");
        }

    const size_t step = 512, sub_elems = (size_t)bInfo.nbricks * step, sub_bytes = sub_elems * sizeof(bElem);
    std::vector<StrongRegion> regions;
    for (size_t i = 0; i < bDecomp.ghost.size(); ++i)
      regions.push_back(StrongRegion{bDecomp.ghost[i].neighbor.set, (size_t)bDecomp.ghost[i].pos * step * 8,
                                     (size_t)bDecomp.skin[i].pos * step * 8, (size_t)bDecomp.ghost[i].len * step * 8});
    StrongPlan sp(dom_size, sdom_size, rk.size, rk.rank, regions);
    const unsigned nsub = (unsigned)(sp.last_id - sp.first_id);

    // two arenas (in / out), subdomain s at base + s * sub_bytes
    char *d_a = nullptr, *d_b = nullptr;
    gpuCheck(gpuMalloc(&d_a, nsub * sub_bytes));
    gpuCheck(gpuMalloc(&d_b, nsub * sub_bytes));
    gpuCheck(bricks_memset(rk.ctx, d_b, 0, nsub * sub_bytes));
    {
      bElem *h = nullptr;
      gpuCheck(bricks_host_alloc((void **)&h, sub_bytes));
      for (unsigned s = 0; s < nsub; ++s) {
#pragma omp parallel for
        for (long l = 0; l < (long)sub_elems; ++l) h[l] = l < 512 ? 0.0 : bricks_detail::uniform01(99 + sp.first_id + s, l);
        gpuCheck(gpuMemcpy(d_a + s * sub_bytes, h, sub_bytes, gpuMemcpyHostToDevice));
      }
      bricks_host_free(h);
    }
    std::vector<const double *> in_a(nsub), in_b(nsub);
    std::vector<double *> out_a(nsub), out_b(nsub);
    for (unsigned s = 0; s < nsub; ++s) {
      in_a[s] = out_a[s] = (double *)(d_a + s * sub_bytes);
      in_b[s] = out_b[s] = (double *)(d_b + s * sub_bytes);
    }

    // link tables and the per-peer messages (strong/main.cu:188-395)
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
    size_t off = 0, tsize = 0;
    for (int r : peers) {
      const StrongPeer *p = find(sp.send, r);
      so.push_back(off), sb.push_back(p ? p->bytes : 0);
      off += (sb.back() + 255) / 256 * 256;
      tsize += sb.back();
    }
    for (int r : peers) {
      const StrongPeer *p = find(sp.recv, r);
      ro.push_back(off), rb.push_back(p ? p->bytes : 0);
      off += (rb.back() + 255) / 256 * 256;
      tsize += rb.back();
    }
    char *arena = nullptr;
    gpuCheck(gpuMalloc(&arena, std::max<size_t>(off, 256)));
    std::vector<bricks_region_link> h_local, h_pack, h_unpack;
    auto link = [](char *from, char *to, size_t len) { return bricks_region_link{(double *)from, (double *)to, len}; };
    for (auto &l : sp.links)
      h_local.push_back(link(d_a + l.src_sub * sub_bytes + l.src_off, d_a + l.dst_sub * sub_bytes + l.dst_off, l.len));
    for (size_t k = 0; k < peers.size(); ++k) {
      if (const StrongPeer *p = find(sp.send, peers[k]))
        for (auto &pc : p->pieces)
          h_pack.push_back(link(d_a + pc.sub * sub_bytes + pc.off, arena + so[k] + pc.buf_off, pc.len));
      if (const StrongPeer *p = find(sp.recv, peers[k]))
        for (auto &pc : p->pieces)
          h_unpack.push_back(link(arena + ro[k] + pc.buf_off, d_a + pc.sub * sub_bytes + pc.off, pc.len));
    }
    Links local, pack, unpack;
    local.upload(rk.ctx, h_local), pack.upload(rk.ctx, h_pack), unpack.upload(rk.ctx, h_unpack);
    bricks_exchange *ex = nullptr;
    if (!peers.empty())
      gpuCheck(bricks_exchange_create_v(rk.ctx, rk.rank, (int)peers.size(), ro.data(), rb.data(), peers.data(), so.data(),
                                        sb.data(), peers.data(), &ex));

    // launch plans over the shared brick grid
    const unsigned gb = GZ / TILE, zero[3] = {0, 0, 0};
    unsigned in_lo[3], in_hi[3], ow_lo[3], ow_hi[3];
    bool has_inner = true;
    for (int a = 0; a < 3; ++a) {
      in_lo[a] = 2 * gb, in_hi[a] = g[a] - 2 * gb, ow_lo[a] = gb, ow_hi[a] = g[a] - gb;
      has_inner = has_inner && in_hi[a] > in_lo[a];
    }
    auto one = [](const unsigned lo[3], const unsigned hi[3]) {
      Box b = {{lo[0], lo[1], lo[2]}, {hi[0], hi[1], hi[2]}};
      return std::vector<Box>{b};
    };
    BrickStencilPlan full(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, one(zero, g));
    BrickStencilPlan owned(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, one(ow_lo, ow_hi));
    std::unique_ptr<BrickStencilPlan> inner, shell;
    if (has_inner && overlap) {
      inner.reset(new BrickStencilPlan(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, one(in_lo, in_hi)));
      shell.reset(new BrickStencilPlan(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, shell_of(g, in_lo, in_hi)));
    }

    const int C_ = BRICKS_STREAM_COMPUTE, X = BRICKS_STREAM_COMM;
    double calltime_l = 0;
    auto exchange = [&]() {
      const double t0 = now();
      gpuCheck(bricks_stream_wait(rk.ctx, X, C_));
      pack.run(X);
      if (ex) gpuCheck(bricks_exchange_run(ex, (double *)arena, X));
      unpack.run(X);
      local.run(X);
      calltime_l += now() - t0;
    };
    auto sweeps = [&](bool naive, bool time_calc) {
      for (int it = 0; it < st_iter; ++it) {
        const auto &src = it % 2 ? in_b : in_a;
        const auto &dst = it % 2 ? out_a : out_b;
        if (naive) {
          gpuCheck(bricks_stencil_naive(full.handle(), which, (int)nsub, src.data(), dst.data(), step, step, nullptr, C_));
        } else if (it == 0) {
          if (inner) {
            if (time_calc) gpuCheck(bricks_timer_start(rk.ctx, C_, 1));
            inner->run(which, src, dst, step, step);
            gpuCheck(bricks_stream_wait(rk.ctx, C_, X));
            shell->run(which, src, dst, step, step);
          } else {
            gpuCheck(bricks_stream_wait(rk.ctx, C_, X));
            if (time_calc) gpuCheck(bricks_timer_start(rk.ctx, C_, 1));
            full.run(which, src, dst, step, step);
          }
        } else if (it == st_iter - 1) {
          owned.run(which, src, dst, step, step);
        } else {
          full.run(which, src, dst, step, step);
        }
      }
      if (time_calc) gpuCheck(bricks_timer_stop(rk.ctx, C_, 1));
    };
    auto brick_func = [&](bool time_calc) {
      exchange();
      sweeps(false, time_calc);
    };

    // ---- validation: same exchange, then the naive kernel's sweeps, vs one real (overlapped) step
    bool ok = true;
    if (validate) {
      char *keep = nullptr, *want = nullptr;
      gpuCheck(gpuMalloc(&keep, nsub * sub_bytes));
      gpuCheck(gpuMalloc(&want, nsub * sub_bytes));
      gpuCheck(gpuMemcpy(keep, d_a, nsub * sub_bytes, gpuMemcpyDeviceToDevice));
      exchange();
      gpuCheck(bricks_stream_wait(rk.ctx, C_, X));
      sweeps(true, false);
      gpuCheck(gpuMemcpy(want, d_a, nsub * sub_bytes, gpuMemcpyDeviceToDevice));
      gpuCheck(gpuMemcpy(d_a, keep, nsub * sub_bytes, gpuMemcpyDeviceToDevice));
      rk.barrier->wait();
      brick_func(false);
      gpuCheck(bricks_sync(rk.ctx, -1));
      std::vector<double> hg(sub_elems), hw(sub_elems);
      double worst = 0;
      for (unsigned s = 0; s < nsub; ++s) {
        gpuCheck(gpuMemcpy(hg.data(), d_a + s * sub_bytes, sub_bytes, gpuMemcpyDeviceToHost));
        gpuCheck(gpuMemcpy(hw.data(), want + s * sub_bytes, sub_bytes, gpuMemcpyDeviceToHost));
        for (unsigned k = gb; k < g[2] - gb; ++k)
          for (unsigned j = gb; j < g[1] - gb; ++j)
            for (unsigned i = gb; i < g[0] - gb; ++i) {
              const size_t b = (size_t)grid_ptr[(k * g[1] + j) * g[0] + i] * step;
              for (unsigned e = 0; e < 512; ++e)
                worst = std::max(worst, std::fabs(hg[b + e] - hw[b + e]) / std::max(std::fabs(hw[b + e]), 1e-300));
            }
      }
      ok = worst <= 1e-12;
      rk.put(9, worst);
      gpuCheck(gpuMemcpy(d_a, keep, nsub * sub_bytes, gpuMemcpyDeviceToDevice));
      gpuFree(keep);
      gpuFree(want);
      rk.barrier->wait();
    }

    // ---- warm-up, timed loop, then a second pass that times the sweeps of each step
    for (int w = 0; w < 3; ++w) brick_func(false);
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    calltime_l = 0;
    gpuCheck(bricks_timer_start(rk.ctx, C_, 0));
    for (int i = 0; i < iters; ++i) brick_func(false);
    gpuCheck(bricks_timer_stop(rk.ctx, C_, 0));
    float ms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 0, &ms));
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    const double call_step = calltime_l / iters;
    double calc_s = 0, pack_s = 0;
    const int cal_iters = std::min(iters, 10);
    for (int i = 0; i < cal_iters; ++i) {
      brick_func(true);
      float cms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 1, &cms));
      calc_s += cms * 1e-3;
    }
    for (int i = 0; i < cal_iters; ++i) {  // pack + unpack + local links alone
      gpuCheck(bricks_timer_start(rk.ctx, X, 2));
      pack.run(X), unpack.run(X), local.run(X);
      gpuCheck(bricks_timer_stop(rk.ctx, X, 2));
      float pms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 2, &pms));
      pack_s += pms * 1e-3;
    }
    const long cnt = (long)iters * st_iter;
    const double total_l = ms * 1e-3 / cnt, calc_l = calc_s / ((double)cal_iters * st_iter);
    const double pack_l = pack_s / ((double)cal_iters * st_iter), exposed = std::max(0.0, total_l - calc_l);
    rk.put(0, total_l);
    rk.put(1, calc_l);
    rk.put(2, pack_l);
    rk.put(3, pack_l > 0 ? tsize / 1.0e9 / (pack_l * st_iter) : 0.0);
    rk.put(4, call_step / st_iter);
    rk.put(5, exposed);
    rk.put(6, (double)tsize * 1.0e-6);
    rk.put(7, exposed > 0 ? tsize / 1.0e9 / (exposed * st_iter) : 0.0);
    rk.put(8, 0.0);
    rk.barrier->wait();
    if (rk.rank == 0) {
      const double total = rk.stats(0).max;
      std::cout << "Bri: " << total << " : " << total << std::endl;
      std::cout << "calc : " << rk.stats(1) << std::endl;
      std::cout << "pack : " << rk.stats(2) << std::endl;
      std::cout << "  | pack speed (GB/s): " << rk.stats(3) << std::endl;
      std::cout << "call : " << rk.stats(4) << std::endl;
      std::cout << "wait : " << rk.stats(5) << std::endl;
      std::cout << "  | MPI size (MB): " << rk.stats(6) << std::endl;
      std::cout << "  | MPI speed (GB/s): " << rk.stats(7) << std::endl;
      std::cout << "move : " << rk.stats(8) << std::endl;
      double perf = dom_size / 1000.0;
      perf = perf * perf * perf / total;
      std::cout << "perf " << perf << " GStencil/s" << std::endl;
      std::cout << "part " << sp.send.size() + sp.recv.size() << std::endl;
    }
    if (validate) {
      if (!ok) {
        std::cout << "result mismatch!" << std::endl;
        std::lock_guard<std::mutex> lk(fail_mutex);
        ++failures;
      }
      rk.barrier->wait();
      if (rk.rank == 0)
        std::cout << (failures ? "Not validated" : "Validated") << ", max relative error vs naive kernel " << rk.stats(9).max
                  << std::endl;
    }
    if (ex) bricks_exchange_destroy(ex);
    gpuFree(arena);
    gpuFree(d_a);
    gpuFree(d_b);
  });
  return rc ? rc : (failures ? 1 : 0);
}
