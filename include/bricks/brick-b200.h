// brick-b200.h -- device glue for host drivers, over the C ABI of libbricks_b200.
//
// Drop-in for the reference's include/brick-cuda.h:12-24 (gpu* macro map) and
// include/brick-gpu.h:23-99 (gpuCheck, movBrickInfo, movBrickStorage).  Host code that includes
// this header never sees cuda_runtime.h; every call goes through include/bricks_b200.h.  Byte
// counts are size_t (the reference keeps them in `unsigned`, which overflows above 4 GiB).
// The macros need a "current" device context: set it once with bricksSetContext(ctx).
#ifndef BRICKS_B200_BRICK_B200_H
#define BRICKS_B200_BRICK_B200_H

#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "../bricks_b200.h"
#include "brick.h"

namespace bricks_detail {
inline bricks_ctx *&current_ctx() {
  static thread_local bricks_ctx *c = nullptr;
  return c;
}
}  // namespace bricks_detail

/// the context the gpu* macros and mov* helpers act on (per host thread)
inline void bricksSetContext(bricks_ctx *ctx) { bricks_detail::current_ctx() = ctx; }
inline bricks_ctx *bricksContext() {
  if (!bricks_detail::current_ctx()) throw std::runtime_error("bricksSetContext() was not called on this thread");
  return bricks_detail::current_ctx();
}

enum gpuMemcpyKind { gpuMemcpyHostToDevice = 1, gpuMemcpyDeviceToHost = 2, gpuMemcpyDeviceToDevice = 3 };
#define gpuSuccess 0
#define gpuGetErrorString(e) bricks_last_error()

inline int gpuMalloc_(void **p, size_t bytes) { return bricks_malloc(bricksContext(), p, bytes); }
#define gpuMalloc(p, s) gpuMalloc_((void **)(p), (size_t)(s))
#define gpuFree(p) bricks_free(bricksContext(), (void *)(p))
inline int gpuMemcpy(void *dst, const void *src, size_t bytes, gpuMemcpyKind kind) {
  bricks_ctx *c = bricksContext();
  int rc = kind == gpuMemcpyHostToDevice   ? bricks_h2d(c, dst, src, bytes, BRICKS_STREAM_COMPUTE)
           : kind == gpuMemcpyDeviceToHost ? bricks_d2h(c, dst, src, bytes, BRICKS_STREAM_COMPUTE)
                                           : bricks_d2d(c, dst, src, bytes, BRICKS_STREAM_COMPUTE);
  return rc ? rc : bricks_sync(c, BRICKS_STREAM_COMPUTE);  // cudaMemcpy semantics: blocking
}
#define gpuDeviceSynchronize() bricks_sync(bricksContext(), -1)

/// Always checks (the reference only checks when NDEBUG is defined, brick-gpu.h:12-20) and
/// throws instead of exit()ing.
#define gpuCheck(x) _gpuCheck((x), #x, __FILE__, __LINE__)
inline void _gpuCheck(int e, const char *func, const char *file, const int line) {
  if (e != gpuSuccess) {
    char msg[1200];
    snprintf(msg, sizeof msg, "\"%s\" at %d in %s\n\treturned %d\n-> %s", func, line, file, e, bricks_last_error());
    throw std::runtime_error(msg);
  }
}

/// Copy the adjacency list to the device (H2D) or back (D2H); the returned BrickInfo's `adj`
/// points at the copy.  Ownership as in the reference: the caller frees it (gpuFree / free).
template <unsigned dims>
BrickInfo<dims> movBrickInfo(BrickInfo<dims> &bInfo, gpuMemcpyKind kind) {
  assert(kind == gpuMemcpyHostToDevice || kind == gpuMemcpyDeviceToHost);
  BrickInfo<dims> ret = bInfo;
  const size_t size = (size_t)bInfo.nbricks * static_power<3, dims>::value * sizeof(unsigned);
  if (kind == gpuMemcpyHostToDevice)
    gpuCheck(gpuMalloc(&ret.adj, size));
  else
    ret.adj = (typename BrickInfo<dims>::adjlist)malloc(size);
  gpuCheck(gpuMemcpy(ret.adj, bInfo.adj, size, kind));
  return ret;
}

/// Copy a whole storage to the device (H2D) or back to fresh host memory (D2H).  The returned
/// storage owns its memory through the shared_ptr deleter (bricks_free / free).
inline BrickStorage movBrickStorage(BrickStorage &bStorage, gpuMemcpyKind kind) {
  assert(kind == gpuMemcpyHostToDevice || kind == gpuMemcpyDeviceToHost);
  const bool to_device = kind == gpuMemcpyHostToDevice;
  BrickStorage ret = bStorage;
  const size_t size = bStorage.step * (size_t)bStorage.chunks * sizeof(bElem);
  bElem *p = nullptr;
  if (to_device)
    gpuCheck(gpuMalloc(&p, size));
  else
    p = (bElem *)malloc(size);
  gpuCheck(gpuMemcpy(p, bStorage.dat.get(), size, kind));
  if (to_device) {
    bricks_ctx *c = bricksContext();
    ret.dat = std::shared_ptr<bElem>(p, [c](bElem *q) { bricks_free(c, q); });
  } else {
    ret.dat = std::shared_ptr<bElem>(p, [](bElem *q) { free(q); });
  }
  return ret;
}

/// copyToBrick / copyFromBrick (bricksetup.h:172-221) between a DEVICE array and a Brick view on a
/// DEVICE storage; grid_dev is the device brick-index grid (BrickStencilPlan::gridDevice()).
/// Same dimlist / padding / ghost meaning as the host versions; asynchronous on `stream`.
template <unsigned dims, typename T>
inline void copyToBrickDevice(const std::vector<long> &dimlist, const std::vector<long> &padding,
                              const std::vector<long> &ghost, bElem *arr_dev, const unsigned *grid_dev, T &brick,
                              int stream = BRICKS_STREAM_COMPUTE) {
  static_assert(dims == 3, "device copies implement the 3-D case");
  static_assert(sizeof(bElem) == 8, "bricks_copy_array moves fp64 elements; fp32 bricks use the host copyToBrick");
  gpuCheck(bricks_copy_array(bricksContext(), 1, dimlist.data(), padding.data(), ghost.data(), (double *)arr_dev, grid_dev,
                             (double *)brick.dat, brick.step, stream));
}
template <unsigned dims, typename T>
inline void copyFromBrickDevice(const std::vector<long> &dimlist, const std::vector<long> &padding,
                                const std::vector<long> &ghost, bElem *arr_dev, const unsigned *grid_dev, T &brick,
                                int stream = BRICKS_STREAM_COMPUTE) {
  static_assert(dims == 3, "device copies implement the 3-D case");
  static_assert(sizeof(bElem) == 8, "bricks_copy_array moves fp64 elements; fp32 bricks use the host copyToBrick");
  gpuCheck(bricks_copy_array(bricksContext(), 0, dimlist.data(), padding.data(), ghost.data(), (double *)arr_dev, grid_dev,
                             (double *)brick.dat, brick.step, stream));
}

namespace bricks_detail {
/// bricks_stencil for the element type of the build: double, or float when the translation unit says
/// `#define bElem float` before including the headers (include/vecscatter.h:12-14 of the reference)
inline int stencil_call(bricks_plan *p, int id, int n, const double *const *in, double *const *out, size_t is, size_t os,
                        const double *c, int stream) {
  return bricks_stencil(p, id, n, in, out, is, os, c, stream);
}
inline int stencil_call(bricks_plan *p, int id, int n, const float *const *in, float *const *out, size_t is, size_t os,
                        const double *c, int stream) {
  return bricks_stencil_f32(p, id, n, in, out, is, os, c, stream);
}
}  // namespace bricks_detail

/// A launch plan over boxes of a brick grid: the host side of
/// `brick_kernel<<<dim3(strideb...), 32>>>` (weak/main.cu:277).  Owns the device copies of the
/// grid and the adjacency list.
class BrickStencilPlan {
 public:
  struct Box {
    unsigned lo[3], hi[3];  // contiguous axis first, hi exclusive
  };
  BrickStencilPlan(bricks_ctx *ctx, const unsigned *grid, const unsigned gdims[3], const unsigned *adj,
                   unsigned nbricks, const std::vector<Box> &boxes) {
    std::vector<unsigned> flat;
    for (const Box &b : boxes) {
      flat.insert(flat.end(), b.lo, b.lo + 3);
      flat.insert(flat.end(), b.hi, b.hi + 3);
    }
    gpuCheck(bricks_plan_create(ctx, grid, gdims, adj, nbricks, (int)boxes.size(), flat.data(), &plan_));
  }
  BrickStencilPlan(const BrickStencilPlan &) = delete;
  BrickStencilPlan &operator=(const BrickStencilPlan &) = delete;
  ~BrickStencilPlan() { bricks_plan_destroy(plan_); }

  /// out = stencil(in) for one pair of Brick views on DEVICE storages (asynchronous)
  template <typename B>
  void run(int stencil_id, B &in, B &out, const double *coeff = nullptr, int stream = BRICKS_STREAM_COMPUTE) {
    const bElem *i[1] = {in.dat};
    bElem *o[1] = {out.dat};
    gpuCheck(bricks_detail::stencil_call(plan_, stencil_id, 1, i, o, in.step, out.step, coeff, stream));
  }
  /// batched over many storage pairs that share the grid/adjacency (strong scaling)
  void run(int stencil_id, const std::vector<const bElem *> &in, const std::vector<bElem *> &out, size_t in_step,
           size_t out_step, const double *coeff = nullptr, int stream = BRICKS_STREAM_COMPUTE) {
    gpuCheck(bricks_detail::stencil_call(plan_, stencil_id, (int)in.size(), in.data(), out.data(), in_step, out_step, coeff,
                                         stream));
  }
  bricks_plan *handle() { return plan_; }
  /// device copies of the grid and the adjacency list held by the plan
  const unsigned *gridDevice() const {
    const unsigned *a = nullptr, *g = nullptr;
    gpuCheck(bricks_plan_device_tables(plan_, &a, &g));
    return g;
  }
  const unsigned *adjDevice() const {
    const unsigned *a = nullptr, *g = nullptr;
    gpuCheck(bricks_plan_device_tables(plan_, &a, &g));
    return a;
  }

 private:
  bricks_plan *plan_ = nullptr;
};

/// The 4-D path: the device side of brick_stencil(out, in, grid_ptr, skip) of weak/main-4d.cpp:62-75 for
/// Brick<Dim<4,4,4,4>, Dim<4>> views (row-major bricks) on DEVICE storages.  Owns the device copies of
/// the grid and the 81-slot adjacency list.
class BrickStencilPlan4D {
 public:
  BrickStencilPlan4D(bricks_ctx *ctx, const unsigned *grid, const unsigned gdims[4], const unsigned *adj,
                     unsigned nbricks) {
    gpuCheck(bricks_plan4d_create(ctx, grid, gdims, adj, nbricks, &plan_));
  }
  BrickStencilPlan4D(const BrickStencilPlan4D &) = delete;
  BrickStencilPlan4D &operator=(const BrickStencilPlan4D &) = delete;
  ~BrickStencilPlan4D() { bricks_plan4d_destroy(plan_); }

  /// out = mpi9pt(in) on every brick at least skip / 4 cells inside the grid (asynchronous); `skip` in
  /// elements, as the reference passes it
  template <typename B>
  void run(B &out, B &in, long skip, int stream = BRICKS_STREAM_COMPUTE) {
    static_assert(B::NDIMS == 4 && B::BRICKSIZE == 256 && B::row_major(), "4x4x4x4 row-major bricks");
    static_assert(sizeof(bElem) == 8, "the 4-D kernels are fp64");
    gpuCheck(bricks_stencil4d(plan_, in.dat, out.dat, in.step, out.step, (unsigned)(skip / 4), stream));
  }
  bricks_plan4d *handle() { return plan_; }

 private:
  bricks_plan4d *plan_ = nullptr;
};

#endif  // BRICKS_B200_BRICK_B200_H
