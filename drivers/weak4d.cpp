// weak4d.cpp -- the 4-D weak-scaling driver: one periodic 4-D subdomain per GPU, 4x4x4x4 bricks, ghost
// depth 8, one halo exchange per ST_ITER = 8 sweeps of the 9-point star of stencils/mpi9pt.py over
// shrinking regions.  Drop-in for the reference's weak/main-4d.cpp (a CPU/MPI program there; flow :77-357)
// with the flags of weak/args.cpp (-d -s -I -b -h), -G ngpus replacing `mpirun -np`, -v for the validation
// against the host ARRAY twin (weak/main-4d.cpp:171-231 array_stencil + exchangeArr, :347-350 compareBrick),
// -a for the page-aligned numbering of the reference's default build (default here: the dense numbering of
// -DDECOMP_PAGEUNALIGN, as in its CUDA drivers) and -t for the verification twin kernel.
//
// Prints the reference's lines (weak/main-4d.cpp:323-343): Bri / calc / call / wait / "  | MPI size (MB)" /
// "  | MPI speed (GB/s)" / "perf X GStencil/s" / "Total of N parts".
#include <unistd.h>

#include <chrono>
#include <cmath>
#include <cstring>

#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"
#include "common.h"
#include "host_twin.h"

namespace {
constexpr unsigned TILE = 4, GZ = 2 * TILE, DIMS = 4;
constexpr int ST_ITER = GZ;  // weak/main-4d.cpp:60
typedef Brick<Dim<4, 4, 4, 4>, Dim<4>> Brick4D;  // fold <4>: row-major bricks, the device layout
typedef BrickDecomp<4, 4, 4, 4, 4> Decomp4D;

const char *help =
    "Running the 4-D weak-scaling brick stencil (d4pt9) on B200s with %s\n\n"
    "Program options\n"
    "  -h: show help (this message)\n"
    "  -G: number of GPUs (ranks), default 1\n"
    "  -b: downsize the process grid to 2-exponential\n"
    "  Domain size, pick either one, in array order contiguous first\n"
    "  -d: comma separated Int[4], overall domain size\n"
    "  -s: comma separated Int[4], per-process domain size\n"
    "  Benchmark control:\n"
    "  -I: number of iterations, default %d\n"
    "  -v: validate against the array computation on the host (periodic exchange between the ranks' arrays)\n"
    "  -a: page-aligned brick numbering (the reference's default build) instead of the dense one\n"
    "  -t: run the verification twin kernel (one thread per point) instead of the marching kernel\n"
    "Example usage:\n"
    "  %s -s 64,64,64,64 -I 20\n";

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

int main(int argc, char **argv) {
  std::vector<unsigned> dom(DIMS, 32);
  int sel = 0, iters = 25, ngpus = 1, c;  // MPI_ITER = 25, weak/main-4d.cpp:78
  bool bin = false, validate = false, aligned = false, twin_kernel = false;
  while ((c = getopt(argc, argv, "d:s:I:hbG:vat")) != -1) {
    switch (c) {
      case 'b': bin = true; break;
      case 'd': drv::parse_tuple(optarg, dom, DIMS); sel = sel ? -2 : 1; break;
      case 's': drv::parse_tuple(optarg, dom, DIMS); sel = sel ? -2 : 2; break;
      case 'I': iters = std::stoi(optarg); break;
      case 'G': ngpus = std::stoi(optarg); break;
      case 'v': validate = true; break;
      case 'a': aligned = true; break;
      case 't': twin_kernel = true; break;
      default: printf("Unknown options %c\n", c);  // fall through
      case 'h': printf(help, argv[0], iters, argv[0]); sel = sel ? -2 : -1;
    }
  }
  if (sel == -2) printf("Contradicting options\n");
  if (sel < 0) return 0;
  if (ngpus < 1 || iters < 1) {
    printf("bad -G / bad -I\n");
    return 1;
  }
  // process grid (weak/args.cpp:92-105)
  int pdims[DIMS] = {1, 1, 1, 1};
  if (bin) {
    int b = 0, s = 1;
    while (s * 2 < ngpus) {
      pdims[b % DIMS] *= 2;
      ++b;
      s *= 2;
    }
  } else {
    cart_dims_create(ngpus, DIMS, pdims);
  }
  const int nranks = bin ? pdims[0] * pdims[1] * pdims[2] * pdims[3] : ngpus;
  const std::vector<unsigned> asked = dom;
  int failures = 0;
  std::mutex fail_mutex;

  int rc = drv::run_ranks(nranks, 8, [&](drv::Rank &rk) {
    BricksComm comm = cart_create(nranks, rk.rank, DIMS, rk.ctx);
    for (unsigned i = 0; i < DIMS; ++i) comm.dims[i] = pdims[i];
    int coo[DIMS];
    cart_coords(comm, rk.rank, coo);

    // split the domain (weak/args.cpp:117-131)
    std::vector<unsigned> dom_size = asked;
    size_t tot_elems;
    if (sel == 1) {
      tot_elems = 1;
      for (unsigned i = 0; i < DIMS; ++i) {
        tot_elems *= dom_size[i];
        unsigned nb = dom_size[i] / TILE, s = nb % pdims[DIMS - 1 - i];
        dom_size[i] = (nb / pdims[DIMS - 1 - i] + (s > (unsigned)coo[DIMS - 1 - i] ? 1 : 0)) * TILE;
      }
    } else {
      tot_elems = nranks;
      for (unsigned i = 0; i < DIMS; ++i) tot_elems *= dom_size[i];
    }
    for (unsigned i = 0; i < DIMS; ++i)
      if (dom_size[i] % TILE || dom_size[i] < 2 * GZ) throw std::runtime_error("subdomain extents must be multiples of 4 and at least 16");
    if (rk.rank == 0) {
      std::cout << "Pagesize " << sysconf(_SC_PAGESIZE) << "; MPI Size " << nranks << " * OpenMP threads 1" << std::endl;
      std::cout << "Domain size of " << tot_elems << " split among" << std::endl;
      std::cout << "A total of " << nranks << " processes " << pdims[0] << "x" << pdims[1] << "x" << pdims[2] << "x" << pdims[3]
                << std::endl;
      std::cout << "d4pt9 on B200, ST_ITER " << ST_ITER << (aligned ? ", page-aligned numbering" : ", dense numbering")
                << (twin_kernel ? ", twin kernel" : "") << std::endl;
    }

    // decomposition (weak/main-4d.cpp:98-140)
    Decomp4D bDecomp(dom_size, GZ);
    bDecomp.comm = comm;
    bDecomp.page_align = aligned;
    populate(comm, bDecomp, 0, 1, coo);
    std::vector<BitSet> skin;
    allneighbors(0, 1, DIMS, skin);
    skin.erase(skin.begin() + (skin.size() / 2));
    bDecomp.initialize(skin);
    BrickInfo<DIMS> bInfo = bDecomp.getBrickInfo();
    const unsigned g[DIMS] = {bDecomp.gridDim(0), bDecomp.gridDim(1), bDecomp.gridDim(2), bDecomp.gridDim(3)};
    const unsigned *grid_ptr = bDecomp.gridData();

    // adjacency symmetry self-check (weak/main-4d.cpp:126-137)
    for (unsigned l = 1; l + 1 < g[3]; ++l)
      for (unsigned k = 1; k + 1 < g[2]; ++k)
        for (unsigned j = 1; j + 1 < g[1]; ++j)
          for (unsigned i = 1; i + 1 < g[0]; ++i) {
            const unsigned b = grid_ptr[((l * g[2] + k) * g[1] + j) * g[0] + i];
            for (unsigned id = 0; id < 81; ++id)
              if (bInfo.adj[bInfo.adj[b][id]][80 - id] != b) throw std::runtime_error("err");
          }

    BrickStorage bStorage = bInfo.allocate(Brick4D::BRICKSIZE);
    BrickStorage bStorageOut = bInfo.allocate(Brick4D::BRICKSIZE);
    memset(bStorage.dat.get(), 0, bStorage.step * bStorage.chunks * sizeof(bElem));  // reserved and padding slots
    memset(bStorageOut.dat.get(), 0, bStorageOut.step * bStorageOut.chunks * sizeof(bElem));
    Brick4D bIn(&bInfo, bStorage, 0), bOut(&bInfo, bStorageOut, 0);
    const std::vector<long> ext = {(long)g[0] * TILE, (long)g[1] * TILE, (long)g[2] * TILE, (long)g[3] * TILE};
    bElem *in_ptr = randomArray(ext, 4321 + rk.rank);
    copyToBrick<DIMS>(ext, in_ptr, const_cast<unsigned *>(grid_ptr), bIn);

    // device side
    BrickStorage bStorage_dev = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
    BrickStorage bStorageOut_dev = movBrickStorage(bStorageOut, gpuMemcpyHostToDevice);
    Brick4D bIn_dev(&bInfo, bStorage_dev, 0), bOut_dev(&bInfo, bStorageOut_dev, 0);
    BrickStencilPlan4D plan(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks);
    ExchangeView ev = bDecomp.exchangeView(bStorage_dev);
    size_t tsize = 0;
    for (auto &gr : bDecomp.ghost) tsize += (size_t)gr.len * bStorage.step * sizeof(bElem) * 2;  // weak/main-4d.cpp:286-288

    auto sweep = [&](Brick4D &out, Brick4D &in, long skip) {
      if (twin_kernel)
        gpuCheck(bricks_stencil4d_naive(plan.handle(), in.dat, out.dat, in.step, out.step, (unsigned)(skip / TILE),
                                        BRICKS_STREAM_COMPUTE));
      else
        plan.run(out, in, skip);
    };
    // the ST_ITER sweeps between two exchanges (weak/main-4d.cpp:266-276)
    auto sweeps = [&]() {
      sweep(bOut_dev, bIn_dev, 0);
      for (int i = 0; i < ST_ITER / 2 - 1; ++i) {
        sweep(bIn_dev, bOut_dev, i * 2 + 1);
        sweep(bOut_dev, bIn_dev, i * 2 + 2);
      }
      sweep(bIn_dev, bOut_dev, GZ);
    };
    double calltime_l = 0;
    auto brick_func = [&]() {
      const double t0 = now();
      ev.exchange();  // on the communication stream; the compute stream waits for it on the device
      sweeps();
      calltime_l += now() - t0;
    };

    bool ok = true;
    if (validate) {
      static std::vector<double *> arrays;  // the ranks' arrays: the host stand-in for exchangeArr
      static std::mutex arrays_mutex;
      {
        std::lock_guard<std::mutex> lk(arrays_mutex);
        arrays.resize(nranks);
        arrays[rk.rank] = in_ptr;
      }
      rk.barrier->wait();
      const long dml[DIMS] = {(long)dom_size[0], (long)dom_size[1], (long)dom_size[2], (long)dom_size[3]};
      twin::fill_ghost4d(in_ptr, arrays, dml, GZ, [&](int sl, int sk, int sj, int si) {
        BitSet n;
        if (si) n.flip(si * 1);
        if (sj) n.flip(sj * 2);
        if (sk) n.flip(sk * 3);
        if (sl) n.flip(sl * 4);
        return bDecomp.rank_map.at(n.set);
      });
      rk.barrier->wait();
      bElem *out_ptr = zeroArray(ext);
      bElem *p = in_ptr, *q = out_ptr;
      for (int it = 0; it < ST_ITER; ++it) {  // the valid region shrinks by one element per sweep
        long lo[DIMS], hi[DIMS];
        for (unsigned a = 0; a < DIMS; ++a) lo[a] = it + 1, hi[a] = ext[a] - it - 1;
        twin::sweep4d(p, q, ext.data(), lo, hi);
        std::swap(p, q);
      }
      brick_func();
      gpuCheck(bricks_sync(rk.ctx, -1));
      BrickStorage got = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      Brick4D bGot(&bInfo, got, 0);
      const std::vector<long> dlist(dml, dml + DIMS), pad0(DIMS, 0), gzv(DIMS, (long)GZ);
      const bool match = compareBrick<DIMS>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bGot);
      const double worst = maxRelErrBrick<DIMS>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bGot);
      ok = match && worst <= 1e-12;
      rk.put(7, worst);
      rk.barrier->wait();
      if (p == out_ptr) std::swap(in_ptr, out_ptr);
      free(out_ptr);
      gpuCheck(gpuMemcpy(bStorage_dev.dat.get(), bStorage.dat.get(), bStorage.step * bStorage.chunks * sizeof(bElem),
                         gpuMemcpyHostToDevice));
    }
    free(in_ptr);

    // warm-up, then the timed loop (stencils/fake.h time_mpi)
    for (int w = 0; w < 3; ++w) brick_func();
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    calltime_l = 0;
    gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    for (int i = 0; i < iters; ++i) brick_func();
    gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    float ms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 0, &ms));
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    // second pass: the sweeps on their own (calc), and one full-grid sweep for the bandwidth line
    const int cal_iters = std::min(iters, 20);
    gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
    for (int i = 0; i < cal_iters; ++i) sweeps();
    gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
    gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 2));
    for (int i = 0; i < cal_iters; ++i) sweep(i % 2 ? bIn_dev : bOut_dev, i % 2 ? bOut_dev : bIn_dev, 0);
    gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 2));
    float cms = 0, fms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 1, &cms));
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 2, &fms));
    const long cnt = (long)iters * ST_ITER;
    const double total_l = ms * 1e-3 / cnt, calc_l = cms * 1e-3 / ((double)cal_iters * ST_ITER);
    rk.put(0, total_l);
    rk.put(1, calc_l);
    rk.put(2, calltime_l / iters / ST_ITER);
    rk.put(3, std::max(0.0, total_l - calc_l));
    rk.put(4, (double)tsize * 1.0e-6);
    rk.put(5, total_l > calc_l ? tsize / 1.0e9 / ((total_l - calc_l) * ST_ITER) : 0.0);
    rk.put(6, fms * 1e-3 / cal_iters);
    rk.barrier->wait();

    if (rk.rank == 0) {
      const mpi_stats tot = rk.stats(0);
      std::cout << "Bri: " << tot.max << std::endl;
      std::cout << "calc " << rk.stats(1) << std::endl;
      std::cout << "call " << rk.stats(2) << std::endl;
      std::cout << "wait " << rk.stats(3) << std::endl;
      std::cout << "  | MPI size (MB): " << rk.stats(4) << std::endl;
      std::cout << "  | MPI speed (GB/s): " << rk.stats(5) << std::endl;
      std::cout << "perf " << (double)tot_elems * 1.0e-9 / tot.max << " GStencil/s" << std::endl;
      std::cout << "Total of " << bDecomp.ghost.size() << " parts" << std::endl;
      const double pts = 256.0 * g[0] * g[1] * g[2] * g[3];
      std::cout << "  | full-grid sweep (ms): " << rk.stats(6).avg * 1e3 << "; HBM (GB/s per GPU, 16 B per point): "
                << 16.0 * pts / rk.stats(6).avg * 1e-9 << std::endl;
    }
    if (validate) {
      if (!ok) {
        std::cout << "result mismatch!" << std::endl;
        std::lock_guard<std::mutex> lk(fail_mutex);
        ++failures;
      }
      rk.barrier->wait();
      if (rk.rank == 0)
        std::cout << (failures ? "result mismatch" : "result match") << ": max relative error vs the host array twin "
                  << rk.stats(7).max << std::endl;
    }
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
  });
  return rc ? rc : (failures ? 1 : 0);
}
