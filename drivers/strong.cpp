// strong.cpp -- strong-scaling driver: a fixed dom^3 periodic grid in (dom/sdom)^3 subdomains named
// by Z-Morton ids, contiguous id ranges per GPU.  Drop-in for the reference's strong/main.cu
// (flow :113-574) and strong/args.cpp (flags -d -s -I -v -h), with -S 7|13|25|125 for the
// stencil, -G ngpus replacing `mpirun -np`, -n for the serial schedule and -v for the validation
// against the GLOBAL periodic array on the host (strong/main.cpp:402-476 validates against the array path).
//
// Per step (strong/main.cu:397-510): pack the skins bound for other GPUs, one message per peer
// (NCCL), unpack into the ghost bricks, device copies between subdomains of the same GPU, then
// ST_ITER sweeps, each ONE launch batched over all local subdomains -- all of it the library's
// bricks_step_run over the program makeStrongStep builds (include/bricks/strong-plan.h), the same
// code bench.py times.  Prints the reference's lines (strong/main.cu:550-566).
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>

#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"
#include "bricks/strong-plan.h"
#include "common.h"
#include "host_twin.h"

namespace {
constexpr unsigned GZ = 8, TILE = 8;
typedef BrickDecomp<3, 8, 8, 8> Decomp3D;
typedef Brick<Dim<8, 8, 8>, Dim<1>> Brick3D;

const char *help =
    "Running the strong-scaling brick stencil on B200s with %s\n\n"
    "Program options\n"
    "  -h: show help (this message)\n"
    "  -G: number of GPUs (ranks), default 1\n"
    "  -d: Int, overall domain size is d^3 (default 512)\n"
    "  -s: Int, subdomain size is s^3 (default 128)\n"
    "  -I: number of iterations, default %d\n"
    "  -S: stencil 7|13|25|125, default 125\n"
    "  -v: validate against the stencil applied to the global periodic array on the host\n"
    "  -n: no overlap (exchange, then sweeps)\n"
    "Example usage:\n"
    "  %s -G 8 -d 1024 -s 128\n";

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

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

    // the step object: box plans over the shared grid + the communication program (pack -> one
    // message per peer -> unpack -> intra-GPU links; strong/main.cu:188-395, 397-447)
    size_t sent = 0;
    bricks_step *stepo = makeStrongStep(rk.ctx, rk.rank, sp, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, GZ / TILE, d_a,
                                        sub_bytes, &sent);
    size_t recvd = 0;
    for (auto &p : sp.recv) recvd += p.bytes;
    const size_t tsize = sent + recvd;
    const int C_ = BRICKS_STREAM_COMPUTE;
    const int flags = overlap ? BRICKS_STEP_OVERLAP : 0;
    double calltime_l = 0;
    auto brick_func = [&]() {
      const double t0 = now();
      gpuCheck(bricks_step_run(stepo, which, st_iter, (int)nsub, out_a.data(), out_b.data(), nullptr, flags));
      calltime_l += now() - t0;
    };

    // ---- validation: one step against the stencil applied to the GLOBAL periodic array on the host
    bool ok = true;
    if (validate) {
      static std::vector<double> glob;  // dom^3, shared by the rank threads
      static std::mutex glob_mutex;
      const long D = dom_size, S = sdom_size, E = S + 2 * GZ;
      {
        std::lock_guard<std::mutex> lk(glob_mutex);
        if (glob.size() != (size_t)D * D * D) glob.assign((size_t)D * D * D, 0.0);
      }
      rk.barrier->wait();
      const std::vector<long> dlist = {S, S, S}, pad0 = {0, 0, 0}, gzv = {(long)GZ, (long)GZ, (long)GZ}, ext = {E, E, E};
      std::vector<double> hs(sub_elems);
      BrickStorage view;
      view.chunks = bInfo.nbricks, view.step = step;
      view.dat = std::shared_ptr<bElem>(hs.data(), [](bElem *) {});
      Brick3D bv(&bInfo, view, 0);
      bElem *arr = zeroArray(ext);
      // every local subdomain's interior into the global array
      for (unsigned s_ = 0; s_ < nsub; ++s_) {
        gpuCheck(gpuMemcpy(hs.data(), d_a + s_ * sub_bytes, sub_bytes, gpuMemcpyDeviceToHost));
        copyFromBrick<3>(dlist, pad0, gzv, arr, const_cast<unsigned *>(grid_ptr), bv);
        unsigned c3[3];
        sp.coords(sp.first_id + s_, c3);
#pragma omp parallel for collapse(2)
        for (long k = 0; k < S; ++k)
          for (long j = 0; j < S; ++j)
            for (long i = 0; i < S; ++i)
              glob[((c3[2] * S + k) * D + c3[1] * S + j) * D + c3[0] * S + i] = arr[((k + GZ) * E + j + GZ) * E + i + GZ];
      }
      rk.barrier->wait();
      brick_func();
      gpuCheck(bricks_sync(rk.ctx, -1));
      double worst = 0;
      bool match = true;
      bElem *tmp = zeroArray(ext);
      const long R = drv::stencil_radius(which);
      for (unsigned s_ = 0; s_ < nsub; ++s_) {
        unsigned c3[3];
        sp.coords(sp.first_id + s_, c3);
        // the subdomain plus its periodic ghost band, cut out of the global array
#pragma omp parallel for collapse(2)
        for (long k = 0; k < E; ++k)
          for (long j = 0; j < E; ++j)
            for (long i = 0; i < E; ++i) {
              const long gk = ((long)c3[2] * S + k - GZ + D) % D, gj = ((long)c3[1] * S + j - GZ + D) % D,
                         gi = ((long)c3[0] * S + i - GZ + D) % D;
              arr[(k * E + j) * E + i] = glob[(gk * D + gj) * D + gi];
            }
        bElem *p = arr, *q = tmp;
        for (int it = 0; it < st_iter; ++it) {
          const long grow = R * (st_iter - 1 - it);
          const long lo[3] = {(long)GZ - grow, (long)GZ - grow, (long)GZ - grow};
          const long hi[3] = {E - GZ + grow, E - GZ + grow, E - GZ + grow};
          twin::sweep(which, p, q, ext.data(), lo, hi);
          std::swap(p, q);
        }
        gpuCheck(gpuMemcpy(hs.data(), d_a + s_ * sub_bytes, sub_bytes, gpuMemcpyDeviceToHost));
        match = compareBrick<3>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bv) && match;
        worst = std::max(worst, maxRelErrBrick<3>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bv));
      }
      free(arr);
      free(tmp);
      ok = match && worst <= 1e-12;
      rk.put(9, worst);
      rk.barrier->wait();
    }

    // ---- warm-up, timed loop, then the sweeps and the exchange program of a step on their own
    for (int w = 0; w < 3; ++w) brick_func();
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    calltime_l = 0;
    gpuCheck(bricks_timer_start(rk.ctx, C_, 0));
    for (int i = 0; i < iters; ++i) brick_func();
    gpuCheck(bricks_timer_stop(rk.ctx, C_, 0));
    float ms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 0, &ms));
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    const double call_step = calltime_l / iters;
    double calc_s = 0, pack_s = 0;
    const int cal_iters = std::min(iters, 10);
    {
      bricks_plan *probe = nullptr;
      const bool split = overlap && bricks_step_plan(stepo, BRICKS_PLAN_INNER, &probe) == 0;
      gpuCheck(bricks_timer_start(rk.ctx, C_, 1));
      for (int i = 0; i < cal_iters; ++i)
        for (int it = 0; it < st_iter; ++it) {
          const auto &src = it % 2 ? in_b : in_a;
          const auto &dst = it % 2 ? out_a : out_b;
          auto sw = [&](int plan) {
            gpuCheck(bricks_step_sweep(stepo, plan, which, (int)nsub, src.data(), dst.data(), nullptr, C_));
          };
          if (it == 0 && split) {
            sw(BRICKS_PLAN_INNER);
            sw(BRICKS_PLAN_SHELL);
          } else if (it == st_iter - 1) {
            sw(BRICKS_PLAN_OWNED);
          } else {
            sw(BRICKS_PLAN_FULL);
          }
        }
      gpuCheck(bricks_timer_stop(rk.ctx, C_, 1));
      float cms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 1, &cms));
      calc_s = cms * 1e-3;
      // the communication program alone (pack, messages, unpack, local links), nothing to hide behind
      gpuCheck(bricks_sync(rk.ctx, -1));
      rk.barrier->wait();
      gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMM, 2));
      for (int i = 0; i < cal_iters; ++i) gpuCheck(bricks_step_exchange(stepo, out_a[0]));
      gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMM, 2));
      float pms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 2, &pms));
      pack_s = pms * 1e-3;
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
      std::cout << "pack : " << rk.stats(2) << std::endl;  // the whole exchange program run alone
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
        std::cout << (failures ? "Not validated" : "Validated") << ", max relative error vs the global periodic array "
                  << rk.stats(9).max << std::endl;
    }
    gpuCheck(bricks_sync(rk.ctx, -1));
    bricks_step_destroy(stepo);
    gpuFree(d_a);
    gpuFree(d_b);
  });
  return rc ? rc : (failures ? 1 : 0);
}
