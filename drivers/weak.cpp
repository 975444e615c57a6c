// weak.cpp -- weak-scaling driver: one periodic subdomain per GPU, one halo exchange per ST_ITER
// stencil sweeps.  Drop-in for the reference's weak/main.cu (flow :45-338) and weak/args.cpp
// (flags -d -s -I -b -h), with -S 7|13|25|125 selecting the stencil at run time (the reference
// picks it with -DMPI_7PT...), -G ngpus replacing `mpirun -np`, -v for the validation against the
// host ARRAY twin (weak/main.cu:143-214, 321-323), -n for the reference's serial schedule (no
// exchange/compute overlap), -E to start every exchange at the step boundary only, -m for the SHIFT
// exchange and -p for one packed message per peer.  The step itself (exchange + ST_ITER sweeps, the
// body of brick_func, weak/main.cu:246-287) is the library's bricks_step_run -- the same code
// bench.py times.
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
#include "host_twin.h"

namespace {
constexpr unsigned GZ = 8, TILE = 8;
typedef Brick<Dim<8, 8, 8>, Dim<1>> Brick3D;
typedef BrickDecomp<3, 8, 8, 8> Decomp3D;

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
    "  -v: validate against the array computation on the host (periodic exchange between the ranks' arrays)\n"
    "  -n: no overlap (exchange, then sweeps); the default on one GPU\n"
    "  -o: overlap on one GPU too\n"
    "  -E: no early exchange (default with more than one GPU: the next step's exchange starts behind the\n"
    "      last sweep's skin bricks)\n"
    "  -m: SHIFT exchange (three stages, six messages) instead of one message per region\n"
    "  -p: one packed message per peer rank (pack kernel, send/recv, unpack kernel; NCCL)\n"
    "  -N: NCCL send/recv for the region messages instead of copy-engine copies into peer-mapped ghost bricks\n"
    "  -e: early exchange also with the copy engines\n"
    "Example usage:\n"
    "  %s -G 8 -s 512,512,512 -S 125\n";

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

int main(int argc, char **argv) {
  std::vector<unsigned> dom(3, 64);
  int sel = 0, iters = 100, ngpus = 1, which = 7, c;
  bool bin = false, validate = false, overlap_flag = true, force_overlap = false, shift = false, packed = false, no_early = false, nccl = false;
  bool force_early = false;
  while ((c = getopt(argc, argv, "d:s:I:hbG:S:vnmpENeo")) != -1) {
    switch (c) {
      case 'b': bin = true; break;
      case 'd': drv::parse_tuple(optarg, dom, 3); sel = sel ? -2 : 1; break;
      case 's': drv::parse_tuple(optarg, dom, 3); sel = sel ? -2 : 2; break;
      case 'I': iters = std::stoi(optarg); break;
      case 'G': ngpus = std::stoi(optarg); break;
      case 'S': which = std::stoi(optarg); break;
      case 'v': validate = true; break;
      case 'n': overlap_flag = false; break;
      case 'o': force_overlap = true; break;
      case 'm': shift = true; break;
      case 'p': packed = true; break;
      case 'E': no_early = true; break;
      case 'e': force_early = true; break;
      case 'N': nccl = true; break;
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
  // one rank: the exchange is an on-device copy of 0.04 ms; hiding it behind a split first sweep costs more than
  // it saves (profiles/r03_scaling/diag_n1), so the serial schedule is the default there (-o forces the overlap)
  const bool overlap = overlap_flag && (ngpus > 1 || force_overlap);

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

    // device side: storages and the library's step object (box plans + communication program)
    BrickStorage bStorage_dev = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
    BrickStorage bStorageOut_dev = movBrickStorage(bStorageOut, gpuMemcpyHostToDevice);
    std::string exchange_text;
    bricks_step *step = bDecomp.makeStep(shift ? Decomp3D::STEP_SHIFT : (packed ? Decomp3D::STEP_PACKED : Decomp3D::STEP_REGIONS),
                                         Brick3D::BRICKSIZE, &exchange_text);
    size_t tsize = 0;
    for (size_t i = 0; i < bDecomp.ghost.size(); ++i)
      tsize += (size_t)(bDecomp.ghost[i].len - bDecomp.ghost[i].first_pad - bDecomp.ghost[i].last_pad) * bStorage.step *
               sizeof(bElem);
    // copy-engine exchange (bricks_step_p2p_*): needs equal subdomains on all ranks (-s, or -d that divides evenly)
    bool p2p = nranks > 1 && !nccl && !packed;
    if (p2p) {
      static std::vector<unsigned char> descs;  // the ranks are threads: "allgather" = a shared table + barrier
      static std::vector<unsigned> doms;
      static std::mutex descs_mutex;
      {
        std::lock_guard<std::mutex> lk(descs_mutex);
        descs.resize((size_t)nranks * 256);
        doms.resize((size_t)nranks * 3);
      }
      rk.barrier->wait();
      for (int a = 0; a < 3; ++a) doms[rk.rank * 3 + a] = dom_size[a];
      gpuCheck(bricks_step_p2p_export(step, bStorage_dev.dat.get(), rk.rank, descs.data() + (size_t)rk.rank * 256));
      rk.barrier->wait();
      bool equal = true;
      for (int r = 0; r < nranks; ++r)
        for (int a = 0; a < 3; ++a) equal = equal && doms[r * 3 + a] == doms[a];
      if (equal)
        gpuCheck(bricks_step_p2p_connect(step, nranks, descs.data()));
      else
        p2p = false;  // unequal split: every rank takes the same decision, the exchange stays on NCCL
      rk.barrier->wait();
    }
    const bool early = overlap && !no_early && nranks > 1 && (force_early || !p2p);
    const int flags = (overlap ? BRICKS_STEP_OVERLAP : 0) | (early ? BRICKS_STEP_EARLY : 0);
    if (rk.rank == 0)
      std::cout << "exchange: " << exchange_text << (p2p ? ", stored straight into peer-mapped ghost bricks" : "")
                << (early ? ", started early behind the last sweep" : "") << std::endl;

    double *fa[1] = {bStorage_dev.dat.get()}, *fb[1] = {bStorageOut_dev.dat.get()};
    double calltime_l = 0;
    // one step = exchange + ST_ITER sweeps (weak/main.cu:246-287); everything is asynchronous
    auto brick_func = [&]() {
      const double t0 = now();
      gpuCheck(bricks_step_run(step, which, st_iter, 1, fa, fb, nullptr, flags));
      calltime_l += now() - t0;
    };

    // validation before the timed loop consumes the initial field: one step against the ARRAY twin
    // on the host (weak/main.cu:143-214 array_stencil with exchangeArr, :321-323 compareBrick)
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
      const long dml[3] = {(long)dom_size[0], (long)dom_size[1], (long)dom_size[2]};
      twin::fill_ghost(in_ptr, arrays, dml, GZ, [&](int sk, int sj, int si) {
        BitSet n;
        if (si) n.flip(si * 1);
        if (sj) n.flip(sj * 2);
        if (sk) n.flip(sk * 3);
        return bDecomp.rank_map.at(n.set);
      });
      rk.barrier->wait();  // nobody sweeps before every ghost band is filled
      bElem *out_ptr = zeroArray(ext);
      bElem *p = in_ptr, *q = out_ptr;
      const long R = drv::stencil_radius(which);
      for (int it = 0; it < st_iter; ++it) {
        const long grow = R * (st_iter - 1 - it);  // the valid region shrinks by R per sweep
        const long lo[3] = {(long)GZ - grow, (long)GZ - grow, (long)GZ - grow};
        const long hi[3] = {ext[0] - GZ + grow, ext[1] - GZ + grow, ext[2] - GZ + grow};
        twin::sweep(which, p, q, ext.data(), lo, hi);
        std::swap(p, q);
      }
      brick_func();
      gpuCheck(bricks_sync(rk.ctx, -1));
      gpuCheck(bricks_step_reset(step));
      BrickStorage got = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      Brick3D bGot(&bInfo, got, 0);
      const std::vector<long> dlist = {dml[0], dml[1], dml[2]}, pad0 = {0, 0, 0}, gzv = {(long)GZ, (long)GZ, (long)GZ};
      const bool match = compareBrick<3>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bGot);
      const double worst = maxRelErrBrick<3>(dlist, pad0, gzv, p, const_cast<unsigned *>(grid_ptr), bGot);
      ok = match && worst <= 1e-12;
      rk.put(7, worst);
      rk.barrier->wait();  // the neighbours have read this rank's array
      if (p == out_ptr) std::swap(in_ptr, out_ptr);  // keep the pair straight for the frees below
      free(out_ptr);
      // restart from the initial field
      gpuCheck(gpuMemcpy(bStorage_dev.dat.get(), bStorage.dat.get(), bStorage.step * bStorage.chunks * sizeof(bElem),
                         gpuMemcpyHostToDevice));
    }
    free(in_ptr);

    // warm-up, then the timed loop (stencils/fake.h time_mpi: barrier, cnt iterations)
    for (int w = 0; w < 3; ++w) brick_func();
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();
    calltime_l = 0;
    gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    for (int i = 0; i < iters; ++i) brick_func();
    gpuCheck(bricks_stream_wait(rk.ctx, BRICKS_STREAM_COMPUTE, BRICKS_STREAM_COMM));  // an early exchange is part of the run
    gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 0));
    float ms = 0;
    gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 0, &ms));
    gpuCheck(bricks_sync(rk.ctx, -1));
    gpuCheck(bricks_step_reset(step));
    rk.barrier->wait();
    const double call_step = calltime_l / iters;
    // second pass: the sweeps of a step on their own (no exchange in flight): calc
    double calc_s = 0;
    const int cal_iters = std::min(iters, 20);
    {
      const double *ia[1] = {fa[0]}, *ib[1] = {fb[0]};
      bricks_plan *probe = nullptr;
      const bool split = overlap && bricks_step_plan(step, BRICKS_PLAN_INNER, &probe) == 0;
      gpuCheck(bricks_timer_start(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
      for (int i = 0; i < cal_iters; ++i)
        for (int it = 0; it < st_iter; ++it) {
          const double *const *src = it % 2 ? ib : ia;
          double *const *dst = it % 2 ? fa : fb;
          auto sw = [&](int plan) {
            gpuCheck(bricks_step_sweep(step, plan, which, 1, src, dst, nullptr, BRICKS_STREAM_COMPUTE));
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
      gpuCheck(bricks_timer_stop(rk.ctx, BRICKS_STREAM_COMPUTE, 1));
      float cms = 0;
      gpuCheck(bricks_timer_elapsed_ms(rk.ctx, 1, &cms));
      calc_s = cms * 1e-3;
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
      if (rk.rank == 0)
        std::cout << (failures ? "result mismatch" : "result match") << ": max relative error vs the host array twin "
                  << rk.stats(7).max << std::endl;
    }
    gpuCheck(bricks_sync(rk.ctx, -1));
    rk.barrier->wait();  // nobody unmaps a storage its peers may still be copying into
    bricks_step_destroy(step);
  });
  return rc ? rc : (failures ? 1 : 0);
}
