// common.h -- what the three drivers share: getopt helpers, the "ranks are host threads" launcher
// that stands in for mpirun, a reusable barrier and the reference's statistics lines.
//
// The reference starts one MPI process per GPU (weak/main.cu:45-60, strong/main.cu:113-128).
// Here one process drives all GPUs of the box: rank r is a host thread that owns device r, its
// own bricks_ctx and its own NCCL communicator rank (-G ngpus replaces `mpirun -np`).
#ifndef BRICKS_B200_DRIVERS_COMMON_H
#define BRICKS_B200_DRIVERS_COMMON_H

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <iostream>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "bricks/brick-b200.h"
#include "bricks/brick-decomp.h"

namespace drv {

/// "a,b,c" -> out[0..dims); a single number fills every axis
inline void parse_tuple(std::string s, std::vector<unsigned> &out, int dims) {
  unsigned last = 0;
  for (int i = 0; i < dims; ++i) {
    if (!s.empty()) {
      size_t r = s.find(',');
      last = (unsigned)std::stoul(s.substr(0, r));
      s = r == std::string::npos ? "" : s.substr(r + 1);
    }
    out[i] = last;
  }
}

/// generation-counting barrier for the rank threads
class Barrier {
 public:
  explicit Barrier(int n) : n_(n) {}
  void wait() {
    std::unique_lock<std::mutex> lk(m_);
    const unsigned long gen = gen_;
    if (++count_ == n_) {
      count_ = 0;
      ++gen_;
      cv_.notify_all();
    } else {
      cv_.wait(lk, [&] { return gen != gen_; });
    }
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  int n_, count_ = 0;
  unsigned long gen_ = 0;
};

/// what every rank thread gets
struct Rank {
  int rank, size;
  bricks_ctx *ctx;
  Barrier *barrier;
  std::vector<std::vector<double>> *slots;  ///< slots[k][rank]: per-rank numbers for mpi_statistics
  void put(int k, double v) { (*slots)[k][rank] = v; }
  mpi_stats stats(int k) const { return mpi_statistics((*slots)[k]); }
};

/// Run body(rank) on `ngpus` threads, one device each; returns non-zero if any rank threw.
inline int run_ranks(int ngpus, int nslots, const std::function<void(Rank &)> &body) {
  int have = 0;
  if (bricks_device_count(&have) != 0 || have < ngpus) {
    fprintf(stderr, "needs %d CUDA device(s), found %d: %s\n", ngpus, have, bricks_last_error());
    return 2;
  }
  unsigned char uid[128];
  if (ngpus > 1 && bricks_comm_unique_id(uid) != 0) {
    fprintf(stderr, "bricks_comm_unique_id: %s\n", bricks_last_error());
    return 2;
  }
  Barrier barrier(ngpus);
  std::vector<std::vector<double>> slots(nslots, std::vector<double>(ngpus, 0.0));
  std::vector<std::string> errors(ngpus);
  std::vector<std::thread> threads;
  for (int r = 0; r < ngpus; ++r)
    threads.emplace_back([&, r] {
      bricks_ctx *ctx = nullptr;
      try {
        gpuCheck(bricks_ctx_create(r, &ctx));
        bricksSetContext(ctx);
        if (ngpus > 1) gpuCheck(bricks_comm_init(ctx, r, ngpus, uid));
        Rank rk{r, ngpus, ctx, &barrier, &slots};
        body(rk);
        gpuCheck(bricks_sync(ctx, -1));
      } catch (const std::exception &e) {
        errors[r] = e.what();
        fprintf(stderr, "rank %d: %s\n", r, e.what());
        // a rank that failed would leave the others waiting in a barrier or a receive
        std::_Exit(3);
      }
      if (ctx) {
        if (ngpus > 1) bricks_comm_destroy(ctx);
        bricks_ctx_destroy(ctx);
      }
    });
  for (auto &t : threads) t.join();
  return 0;
}

inline int stencil_radius(int which) { return which == 7 ? 1 : which == 25 ? 4 : 2; }
/// sweeps per exchange with an 8-deep ghost zone (stencils/fake.h: ST_ITER = GZ / radius)
inline int stencil_iters(int which) { return 8 / stencil_radius(which); }
inline bool stencil_known(int which) { return which == 7 || which == 13 || which == 25 || which == 125; }

}  // namespace drv

#endif  // BRICKS_B200_DRIVERS_COMMON_H
