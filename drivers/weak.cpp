// weak.cpp -- weak-scaling driver: one periodic subdomain per GPU, one halo exchange per ST_ITER
// stencil sweeps.  Drop-in for the reference's weak/main.cu (flow :45-338) and weak/args.cpp
// (flags -d -s -I -b -h), with -S 7|13|25|125 selecting the stencil at run time (the reference
// picks it with -DMPI_7PT...), -G ngpus replacing `mpirun -np`, -v for the validation twin and
// -n for the reference's serial schedule (no exchange/compute overlap).  -m uses the SHIFT exchange.
//
// Prints the reference's lines (weak/main.cu:304-318): Bri / calc / move / call / wait /
// "  | MPI size (MB)" / "  | MPI speed (GB/s)" / "perf X GStencil/s".
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstring>

#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"
#include "common.h"

namespace {
constexpr unsigned GZ = 8, TILE = 8;
typedef Brick<Dim<8, 8, 8>, Dim<1>> Brick3D;
typedef BrickDecomp<3, 8, 8, 8> Decomp3D;
typedef BrickStencilPlan::Box Box;

const char *help =
    "Running the weak-scaling brick stencil on B200s with %s\n\n"
    "Program options\n"
    "  -h: show help (this message)\n"
    "  -G: number of GPUs (ranks), default 1\n"
    "  -b: downsize the process grid to 2-exponential\n"
    "  Domain size, pick either one, in array order contiguous first\n"
    "  -d: comma separated Int[3], overall domain size\n"
    "  -s: comma separated Int[3], per-process domain size\n"
    "  Benchmark control:\n"
    "  -I: number of iterations, default %d\n"
    "  -S: stencil 7|13|25|125, default 7\n"
    "  -v: validate against the naive device kernel\n"
    "  -n: no overlap (exchange, then sweeps)\n"
    "  -m: SHIFT exchange (three stages, six messages) instead of one message per region\n"
    "Example usage:\n"
    "  %s -G 8 -s 512,512,512 -S 125\n";

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

/// six slabs covering [0,g) minus [lo,hi)
std::vector<Box> shell_of(const unsigned g[3], const unsigned lo[3], const unsigned hi[3]) {
  auto box = [](unsigned a0, unsigned a1, unsigned a2, unsigned b0, unsigned b1, unsigned b2) {
    Box b = {{a0, a1, a2}, {b0, b1, b2}};
    return b;
  };
  return {box(0, 0, 0, g[0], g[1], lo[2]),         box(0, 0, hi[2], g[0], g[1], g[2]),
          box(0, 0, lo[2], g[0], lo[1], hi[2]),    box(0, hi[1], lo[2], g[0], g[1], hi[2]),
          box(0, lo[1], lo[2], lo[0], hi[1], hi[2]), box(hi[0], lo[1], lo[2], g[0], hi[1], hi[2])};
}
}  // namespace

int main(int argc, char **argv) {
  std::vector<unsigned> dom(3, 64);
  int sel = 0, iters = 100, ngpus = 1, which = 7, c;
  bool bin = false, validate = false, overlap = true, shift = false;
  while ((c = getopt(argc, argv, "d:s:I:hbG:S:vnm")) != -1) {
    switch (c) {
      case 'b': bin = true; break;
      case 'd': drv::parse_tuple(optarg, dom, 3); sel = sel ? -2 : 1; break;
      case 's': drv::parse_tuple(optarg, dom, 3); sel = sel ? -2 : 2; break;
      case 'I': iters = std::stoi(optarg); break;
      case 'G': ngpus = std::stoi(optarg); break;
      case 'S': which = std::stoi(optarg); break;
      case 'v': validate = true; break;
      case 'n': overlap = false; break;
      case 'm': shift = true; break;
      default: printf("Unknown options %c\n", c);  // fall through
      case 'h': printf(help, argv[0], iters, argv[0]); sel = sel ? -2 : -1;
    }
  }
  if (sel == -2) printf("Contradicting options\n");
  if (sel < 0) return 0;
  if (!drv::stencil_known(which) || ngpus < 1 || iters < 1) {
    printf("Unknown stencil / bad -G / bad -I\n");
    return 1;
  }
  const int st_iter = drv::stencil_iters(which);

  // process grid (weak/args.cpp:92-105)
  int pdims[3] = {1, 1, 1};
  if (bin) {
    int b = 0, s = 1;
    while (s * 2 < ngpus) {
      pdims[b % 3] *= 2;
      ++b;
      s *= 2;
    }
  } else {
    cart_dims_create(ngpus, 3, pdims);
  }
  const int nranks = bin ? pdims[0] * pdims[1] * pdims[2] : ngpus;
  const std::vector<unsigned> asked = dom;
  int failures = 0;
  std::mutex fail_mutex;

  int rc = drv::run_ranks(nranks, 8, [&](drv::Rank &rk) {
    BricksComm comm = cart_create(nranks, rk.rank, 3, rk.ctx);
    for (int i = 0; i < 3; ++i) comm.dims[i] = pdims[i];
    int coo[3];
    cart_coords(comm, rk.rank, coo);

    // split the domain (weak/args.cpp:117-131)
    std::vector<unsigned> dom_size = asked;
    size_t tot_elems;
    if (sel == 1) {
      tot_elems = 1;
      for (int i = 0; i < 3; ++i) {
        tot_elems *= dom_size[i];
        unsigned nb = dom_size[i] / TILE, s = nb % pdims[2 - i];
        dom_size[i] = (nb / pdims[2 - i] + (s > (unsigned)coo[2 - i] ? 1 : 0)) * TILE;
      }
    } else {
      tot_elems = nranks;
      for (int i = 0; i < 3; ++i) tot_elems *= dom_size[i];
    }
    if (rk.rank == 0) {
      std::cout << "Pagesize " << sysconf(_SC_PAGESIZE) << "; MPI Size " << nranks << " * OpenMP threads 1" << std::endl;
      std::cout << "Domain size of " << tot_elems << " split among" << std::endl;
      std::cout << "A total of " << nranks << " processes " << pdims[0] << "x" << pdims[1] << "x" << pdims[2] << std::endl;
      std::cout << "d3pt" << which << " on B200, ST_ITER " << st_iter << (overlap ? ", exchange overlapped" : ", serial")
                << std::endl;
    }

    // decomposition (weak/main.cu:63-100)
    Decomp3D bDecomp(dom_size, GZ);
    bDecomp.comm = comm;
    populate(comm, bDecomp, 0, 1, coo);
    bDecomp.initialize(skin3d_good);
    BrickInfo<3> bInfo = bDecomp.getBrickInfo();
    const unsigned g[3] = {bDecomp.gridDim(0), bDecomp.gridDim(1), bDecomp.gridDim(2)};
    const unsigned *grid_ptr = bDecomp.gridData();

    // adjacency symmetry self-check (weak/main.cu:102-109)
    for (unsigned k = 1; k + 1 < g[2]; ++k)
      for (unsigned j = 1; j + 1 < g[1]; ++j)
        for (unsigned i = 1; i + 1 < g[0]; ++i) {
          const unsigned l = grid_ptr[(k * g[1] + j) * g[0] + i];
          for (unsigned id = 0; id < 27; ++id)
            if (bInfo.adj[bInfo.adj[l][id]][26 - id] != l) throw std::runtime_error("err");
        }

    BrickStorage bStorage = bInfo.allocate(Brick3D::BRICKSIZE);
    BrickStorage bStorageOut = bInfo.allocate(Brick3D::BRICKSIZE);
    Brick3D bIn(&bInfo, bStorage, 0), bOut(&bInfo, bStorageOut, 0);
    const std::vector<long> ext = {(long)g[0] * TILE, (long)g[1] * TILE, (long)g[2] * TILE};
    bElem *in_ptr = randomArray(ext, 1234 + rk.rank);
    copyToBrick<3>(ext, in_ptr, const_cast<unsigned *>(grid_ptr), bIn);
    memset(bStorageOut.dat.get(), 0, bStorageOut.step * bStorageOut.chunks * sizeof(bElem));
    free(in_ptr);

    // device side: storages, launch plans, exchange view
    BrickStorage bStorage_dev = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
    BrickStorage bStorageOut_dev = movBrickStorage(bStorageOut, gpuMemcpyHostToDevice);
    Brick3D bIn_dev(&bInfo, bStorage_dev, 0), bOut_dev(&bInfo, bStorageOut_dev, 0);
    const unsigned gb = GZ / TILE;
    const unsigned zero[3] = {0, 0, 0};
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
    ExchangeView ev_direct = bDecomp.exchangeView(bStorage_dev);
    size_t tsize = 0;
    for (size_t n : ev_direct.seclen) tsize += n;
    // -m: the SHIFT exchange (multiStageExchangeView, brick-mpi.h:587-681): 3 stages, 6 messages
    MultiStageExchangeView ev_shift;
    if (shift) ev_shift = bDecomp.multiStageExchangeView(bStorage_dev);
    struct {
      ExchangeView *d;
      MultiStageExchangeView *m;
      void start() { m ? m->start() : d->start(); }
      void wait() { m ? m->wait() : d->wait(); }
    } ev = {&ev_direct, shift ? &ev_shift : nullptr};

    double calltime_l = 0;
    // one step = exchange + ST_ITER sweeps (weak/main.cu:246-287); everything is asynchronous
    auto brick_func = [&](bool time_calc) {
      const double t0 = now();
      ev.start();
      calltime_l += now() - t0;
      Brick3D *src = &bIn_dev, *dst = &bOut_dev;
      for (int it = 0; it < st_iter; ++it) {
        if (it == 0) {
          if (inner) {
            if (time_calc) gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
            inner->run(which, *src, *dst);
            ev.wait();
            shell->run(which, *src, *dst);
          } else {
            ev.wait();
            if (time_calc) gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
            full.run(which, *src, *dst);
          }
        } else if (it == st_iter - 1) {
          owned.run(which, *src, *dst);
        } else {
          full.run(which, *src, *dst);
        }
        std::swap(src, dst);
      }
      if (time_calc) gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
    };

    // validation twin before the timed loop consumes the initial field
    bool ok = true;
    if (validate) {
      BrickStorage a2 = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
      BrickStorage b2 = movBrickStorage(bStorageOut, gpuMemcpyHostToDevice);
      bDecomp.exchange(a2);
      const double *i2[1];
      double *o2[1];
      bElem *p = a2.dat.get(), *q = b2.dat.get();
      for (int it = 0; it < st_iter; ++it) {
        i2[0] = p, o2[0] = q;
        gpuCheck(bricks_stencil_naive(full.handle(), which, 1, i2, o2, a2.step, b2.step, nullptr, BRICKS_STREAM_COMPUTE));
        std::swap(p, q);
      }
      rk.barrier->wait();
      brick_func(false);
      gpuCheck(bricks_sync(rk.ctx, -1));
      BrickStorage got = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      BrickStorage want = movBrickStorage(a2, gpuMemcpyDeviceToHost);
      double worst = 0;
      for (unsigned k = gb; k < g[2] - gb; ++k)
        for (unsigned j = gb; j < g[1] - gb; ++j)
          for (unsigned i = gb; i < g[0] - gb; ++i) {
            const size_t b = (size_t)grid_ptr[(k * g[1] + j) * g[0] + i] * got.step;
            for (unsigned e = 0; e < 512; ++e) {
              const double w = want.dat.get()[b + e], d = std::fabs(got.dat.get()[b + e] - w);
              worst = std::max(worst, d / std::max(std::fabs(w), 1e-300));
            }
          }
      ok = worst <= 1e-12;
      rk.put(7, worst);
      // restart from the initial field
      gpuCheck(gpuMemcpy(bStorage_dev.dat.get(), bStorage.dat.get(), bStorage.step * bStorage.chunks * sizeof(bElem),
                         gpuMemcpyHostToDevice));
    }

    // warm-up, then the timed loop (stencils/fake.h time_mpi: barrier, cnt iterations)
    for (int w = 0; w < 3; ++w) brick_func(false);
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    calltime_l = 0;
    gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    for (int i = 0; i < iters; ++i) brick_func(false);
    gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    float ms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 0, &ms));
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    const double call_step = calltime_l / iters;
    // second pass: time the sweeps of each step on their own (the reference synchronises per step too)
    double calc_s = 0;
    const int cal_iters = std::min(iters, 20);
    for (int i = 0; i < cal_iters; ++i) {
      brick_func(true);
      float cms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 1, &cms));
      calc_s += cms * 1e-3;
    }
    const long cnt = (long)iters * st_iter;
    const double total_l = ms * 1e-3 / cnt, calc_l = calc_s / ((double)cal_iters * st_iter);
    rk.put(0, total_l);
    rk.put(1, calc_l);
    rk.put(2, call_step / st_iter);
    rk.put(3, std::max(0.0, total_l - calc_l));
    rk.put(4, (double)tsize * 1.0e-6);
    rk.put(5, std::max(0.0, total_l - calc_l) > 0 ? tsize / 1.0e9 / (std::max(1e-9, total_l - calc_l) * st_iter) : 0.0);
    rk.put(6, 0.0);
    rk.barrier->wait();

    if (rk.rank == 0) {
      const mpi_stats tot = rk.stats(0);
      // the reference prints the SUM of the phase averages (no overlap); with overlap the step
      // time is the max over ranks of the measured wall time per sweep
      std::cout << "Bri: " << tot.max << std::endl;
      std::cout << "calc " << rk.stats(1) << std::endl;
      std::cout << "move " << rk.stats(6) << std::endl;
      std::cout << "call " << rk.stats(2) << std::endl;
      std::cout << "wait " << rk.stats(3) << std::endl;
      std::cout << "  | MPI size (MB): " << rk.stats(4) << std::endl;
      std::cout << "  | MPI speed (GB/s): " << rk.stats(5) << std::endl;
      std::cout << "perf " << (double)tot_elems * 1.0e-9 / tot.max << " GStencil/s" << std::endl;
      const double bytes = 16.0 * g[0] * g[1] * g[2] * 512;
      std::cout << "  | HBM (GB/s per GPU, full sweeps): " << bytes / rk.stats(1).avg * 1e-9 << std::endl;
    }
    if (validate) {
      if (!ok) {
        std::cout << "result mismatch!" << std::endl;
        std::lock_guard<std::mutex> lk(fail_mutex);
        ++failures;
      }
      rk.barrier->wait();
      if (rk.rank == 0) std::cout << "max relative error vs naive kernel " << rk.stats(7).max << std::endl;
    }
  });
  return rc ? rc : (failures ? 1 : 0);
}
