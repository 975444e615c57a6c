// single.cpp -- single-GPU driver.  Drop-in for the reference's single/cpu.cpp:10-25 and
// single/cuda.cpp:13-24 (d3pt7 / d3pt7cu of stencils/3axis.cpp:14-83 and 3axis.cu:84-190):
// 129 random coefficients, an N^3 grid with one ghost brick per side and PADDING elements of
// array padding, the interleaved in/out storage (step = 2 * brick size), the CPU array stencil
// as "Arr" and the B200 brick kernel as "Bri", then compareBrick.  After the reference's d3pt7
// test it runs the fixed-coefficient 7/13/25/125-point stencils (stencils/fake.h MPI_*PT) on the
// same grid and checks them against the naive device kernel.
//
//   single [-d N] [-S 7|13|25|125|0 (0 = all)] [-I launches per timing]
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>

#include "bricks/brick-b200.h"
#include "bricks/brickcompare.h"
#include "bricks/bricksetup.h"
#include "bricks/multiarray.h"
#include "common.h"

namespace {
constexpr long TILE = 8, GZ = 8, PADDING = 8;
typedef Brick<Dim<8, 8, 8>, Dim<1>> Brick3D;
bElem coeff[129];

/// the reference's cutime_func (stencils/stencils_cu.h:13-35): one warm-up call, then average
/// seconds per call over `reps` back-to-back calls between two events
template <typename F>
double cutime_func(bricks_ctx *ctx, F func, int reps) {
  func();
  gpuCheck(bricks_timer_start(ctx, BRICKS_STREAM_COMPUTE, 0));
  for (int i = 0; i < reps; ++i) func();
  gpuCheck(bricks_timer_stop(ctx, BRICKS_STREAM_COMPUTE, 0));
  float ms = 0;
  gpuCheck(bricks_timer_elapsed_ms(ctx, 0, &ms));
  return ms * 1e-3 / reps;
}

template <typename F>
double time_host(F func, int reps) {
  func();
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; ++i) func();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / reps;
}
}  // namespace

int main(int argc, char **argv) {
  long N = 512;
  int which = 0, reps = 20, c;
  while ((c = getopt(argc, argv, "d:S:I:h")) != -1) {
    switch (c) {
      case 'd': N = std::stol(optarg); break;
      case 'S': which = std::stoi(optarg); break;
      case 'I': reps = std::stoi(optarg); break;
      default:
        printf("usage: %s [-d N] [-S 0|1|7|13|25|27|125] [-I reps]\n", argv[0]);
        return c == 'h' ? 0 : 1;
    }
  }
  if (N % TILE || N < 2 * TILE || (which && which != 1 && which != 27 && !drv::stencil_known(which))) {
    printf("N must be a multiple of %ld (>= %ld); stencil one of 1 (cond), 7, 13, 25, 27, 125\n", TILE, 2 * TILE);
    return 1;
  }
  for (int i = 0; i < 129; ++i) coeff[i] = bricks_detail::uniform01(42 ^ 0x5EED, i);

  int failures = 0;
  int rc = drv::run_ranks(1, 1, [&](drv::Rank &rk) {
    const long STRIDE = N + 2 * (GZ + PADDING), STRIDEG = N + 2 * GZ, STRIDEB = STRIDEG / TILE, NB = N / TILE;
    // grid + adjacency: init_grid over the STRIDEB^3 brick grid (stencils/3axis.cpp:15-20)
    unsigned *grid_ptr;
    BrickInfo<3> bInfo = init_grid<3>(grid_ptr, {STRIDEB, STRIDEB, STRIDEB});
    bElem *in_ptr = randomArray({STRIDE, STRIDE, STRIDE}, 7);
    bElem *out_ptr = zeroArray({STRIDE, STRIDE, STRIDE});

    const size_t bSize = Brick3D::BRICKSIZE;
    BrickStorage bStorage = BrickStorage::allocate(bInfo.nbricks, bSize * 2);  // in and out interleaved
    memset(bStorage.dat.get(), 0, bStorage.step * bStorage.chunks * sizeof(bElem));
    Brick3D bIn(&bInfo, bStorage, 0), bOut(&bInfo, bStorage, bSize);
    copyToBrick<3>({STRIDEG, STRIDEG, STRIDEG}, {PADDING, PADDING, PADDING}, {0, 0, 0}, in_ptr, grid_ptr, bIn);
    BrickStorage bStorage_dev = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
    Brick3D bIn_dev(&bInfo, bStorage_dev, 0), bOut_dev(&bInfo, bStorage_dev, bSize);

    // launch over the NB^3 interior bricks: dim3 block(NB, NB, NB) at offset GB (3axis.cu:28-37)
    const unsigned g[3] = {(unsigned)STRIDEB, (unsigned)STRIDEB, (unsigned)STRIDEB};
    BrickStencilPlan::Box box = {{1, 1, 1}, {(unsigned)(1 + NB), (unsigned)(1 + NB), (unsigned)(1 + NB)}};
    BrickStencilPlan plan(rk.ctx, grid_ptr, g, &bInfo.adj[0][0], bInfo.nbricks, {box});
    const double pts = (double)N * N * N;

    auto check = [&](const char *name, BrickStorage &got, bElem *want_arr) {
      Brick3D b(&bInfo, got, bSize);
      if (!compareBrick<3>({N, N, N}, {PADDING, PADDING, PADDING}, {GZ, GZ, GZ}, want_arr, grid_ptr, b)) {
        std::cout << name << ": result mismatch!" << std::endl;
        ++failures;
      }
    };

    if (which == 0 || which == 7) {
      // ---- the reference's d3pt7 test: random coefficients, CPU array twin
      auto arr_func = [&]() {
        const long s1 = STRIDE, s2 = STRIDE * STRIDE, o = GZ + PADDING;
#pragma omp parallel for collapse(2)
        for (long k = o; k < o + N; ++k)
          for (long j = o; j < o + N; ++j) {
            const bElem *a = in_ptr + k * s2 + j * s1;
            bElem *r = out_ptr + k * s2 + j * s1;
#pragma omp simd
            for (long i = o; i < o + N; ++i)
              r[i] = coeff[5] * a[i + s2] + coeff[6] * a[i - s2] + coeff[3] * a[i + s1] + coeff[4] * a[i - s1] +
                     coeff[1] * a[i + 1] + coeff[2] * a[i - 1] + coeff[0] * a[i];
          }
      };
      auto brick_func = [&]() { plan.run(BRICKS_ST_7PT_COEFF, bIn_dev, bOut_dev, coeff); };
      std::cout << "d3pt7" << std::endl;
      const double t_arr = time_host(arr_func, 3);
      std::cout << "Arr: " << t_arr << std::endl;
      const double t_bri = cutime_func(rk.ctx, brick_func, reps);
      std::cout << "Bri: " << t_bri << std::endl;
      std::cout << "  | Arr (host OpenMP) " << pts / t_arr * 1e-9 << " GStencil/s, Bri (B200) " << pts / t_bri * 1e-9
                << " GStencil/s, " << 16.0 * pts / t_bri * 1e-9 << " GB/s" << std::endl;
      BrickStorage back = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      check("d3pt7", back, out_ptr);
      // ---- the array layout on the device: the "Arr" line of d3pt7cu (stencils/3axis.cu:39-60,173-174),
      // compared with the host array result element by element
      {
        const size_t abytes = (size_t)STRIDE * STRIDE * STRIDE * sizeof(bElem);
        bElem *ain_dev = nullptr, *aout_dev = nullptr;
        gpuCheck(gpuMalloc(&ain_dev, abytes));
        gpuCheck(gpuMalloc(&aout_dev, abytes));
        gpuCheck(gpuMemcpy(ain_dev, in_ptr, abytes, gpuMemcpyHostToDevice));
        gpuCheck(bricks_memset(rk.ctx, aout_dev, 0, abytes));
        const long ext[3] = {STRIDE, STRIDE, STRIDE}, o = GZ + PADDING, lo[3] = {o, o, o}, hi[3] = {o + N, o + N, o + N};
        auto cuarr_func = [&]() {
          gpuCheck(bricks_array_stencil7(rk.ctx, ain_dev, aout_dev, ext, lo, hi, coeff, BRICKS_STREAM_COMPUTE));
        };
        const double t_cuarr = cutime_func(rk.ctx, cuarr_func, reps);
        std::cout << "Arr (B200): " << t_cuarr << std::endl;
        std::cout << "  | " << pts / t_cuarr * 1e-9 << " GStencil/s, " << 16.0 * pts / t_cuarr * 1e-9
                  << " GB/s; brick layout / array layout = " << t_cuarr / t_bri << "x" << std::endl;
        std::vector<bElem> got((size_t)STRIDE * STRIDE * STRIDE);
        gpuCheck(gpuMemcpy(got.data(), aout_dev, abytes, gpuMemcpyDeviceToHost));
        if (!compareArray({STRIDE, STRIDE, STRIDE}, got.data(), out_ptr)) {  // both are zero outside the box
          std::cout << "d3pt7 array kernel: result mismatch!" << std::endl;
          ++failures;
        }
        gpuCheck(gpuFree(ain_dev));
        gpuCheck(gpuFree(aout_dev));
      }
    }

    if (which == 0 || which == 1) {
      // ---- the reference's d3cond / d3condcu test (stencils/3axis.cpp:94-163, cond.py): max(in, 0) per
      // term, |sum| as the result; on the reference's U[0,1) field and coefficients both are no-ops
      auto arr_func = [&]() {
        const long s1 = STRIDE, s2 = STRIDE * STRIDE, o = GZ + PADDING;
#pragma omp parallel for collapse(2)
        for (long k = o; k < o + N; ++k)
          for (long j = o; j < o + N; ++j) {
            const bElem *a = in_ptr + k * s2 + j * s1;
            bElem *r = out_ptr + k * s2 + j * s1;
#pragma omp simd
            for (long i = o; i < o + N; ++i) {
              const bElem partial = coeff[5] * std::max(a[i + s2], 0.0) + coeff[6] * std::max(a[i - s2], 0.0) +
                                    coeff[3] * std::max(a[i + s1], 0.0) + coeff[4] * std::max(a[i - s1], 0.0) +
                                    coeff[1] * std::max(a[i + 1], 0.0) + coeff[2] * std::max(a[i - 1], 0.0) +
                                    coeff[0] * std::max(a[i], 0.0);
              r[i] = partial > 0 ? partial : -partial;
            }
          }
      };
      auto brick_func = [&]() { plan.run(BRICKS_ST_COND, bIn_dev, bOut_dev, coeff); };
      std::cout << "cond" << std::endl;
      const double t_arr = time_host(arr_func, 3);
      std::cout << "Arr: " << t_arr << std::endl;
      const double t_bri = cutime_func(rk.ctx, brick_func, reps);
      std::cout << "Bri: " << t_bri << std::endl;
      std::cout << "  | Arr (host OpenMP) " << pts / t_arr * 1e-9 << " GStencil/s, Bri (B200) " << pts / t_bri * 1e-9
                << " GStencil/s, " << 16.0 * pts / t_bri * 1e-9 << " GB/s" << std::endl;
      BrickStorage back = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      check("cond", back, out_ptr);
    }

    if (which == 0 || which == 27) {
      // ---- the reference's d3pt27 test (stencils/3axis.cpp:165-247): 27 coefficients; term 16 is read
      // at [k+1][j+1][i] in both of its forms, and so it is here
      auto arr_func = [&]() {
        const long s1 = STRIDE, s2 = STRIDE * STRIDE, o = GZ + PADDING;
#pragma omp parallel for collapse(2)
        for (long k = o; k < o + N; ++k)
          for (long j = o; j < o + N; ++j)
            for (long i = o; i < o + N; ++i) {
              const bElem *a = in_ptr + k * s2 + j * s1 + i;
              bElem r = 0;
              for (int t = 0; t < 27; ++t) {
                const long dk = t == 16 ? 1 : t / 9 - 1, dj = (t / 3) % 3 - 1, di = t % 3 - 1;
                r = t ? r + coeff[t] * a[dk * s2 + dj * s1 + di] : coeff[t] * a[dk * s2 + dj * s1 + di];
              }
              out_ptr[k * s2 + j * s1 + i] = r;
            }
      };
      auto brick_func = [&]() { plan.run(BRICKS_ST_27PT_COEFF, bIn_dev, bOut_dev, coeff); };
      std::cout << "d3pt27" << std::endl;
      const double t_arr = time_host(arr_func, 3);
      std::cout << "Arr: " << t_arr << std::endl;
      const double t_bri = cutime_func(rk.ctx, brick_func, reps);
      std::cout << "Bri: " << t_bri << std::endl;
      std::cout << "  | Arr (host OpenMP) " << pts / t_arr * 1e-9 << " GStencil/s, Bri (B200) " << pts / t_bri * 1e-9
                << " GStencil/s, " << 16.0 * pts / t_bri * 1e-9 << " GB/s" << std::endl;
      BrickStorage back = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      check("d3pt27", back, out_ptr);
    }

    // ---- fixed-coefficient stencils of the MPI drivers on the same grid, vs the naive kernel
    for (int w : {7, 13, 25, 125}) {
      if (which && which != w) continue;  // -S 1 / -S 27 run cond / d3pt27 only
      BrickStorage twin = movBrickStorage(bStorage, gpuMemcpyHostToDevice);
      const double *i2[1] = {twin.dat.get()};
      double *o2[1] = {twin.dat.get() + bSize};
      gpuCheck(bricks_stencil_naive(plan.handle(), w, 1, i2, o2, twin.step, twin.step, nullptr, BRICKS_STREAM_COMPUTE));
      auto brick_func = [&]() { plan.run(w, bIn_dev, bOut_dev); };
      std::cout << "mpi" << w << "pt" << std::endl;
      const double t_bri = cutime_func(rk.ctx, brick_func, reps);
      std::cout << "Bri: " << t_bri << std::endl;
      std::cout << "  | " << pts / t_bri * 1e-9 << " GStencil/s, " << 16.0 * pts / t_bri * 1e-9 << " GB/s" << std::endl;
      BrickStorage got = movBrickStorage(bStorage_dev, gpuMemcpyDeviceToHost);
      BrickStorage want = movBrickStorage(twin, gpuMemcpyDeviceToHost);
      double worst = 0;
      for (long k = 1; k <= NB; ++k)
        for (long j = 1; j <= NB; ++j)
          for (long i = 1; i <= NB; ++i) {
            const size_t b = (size_t)grid_ptr[(k * STRIDEB + j) * STRIDEB + i] * got.step + bSize;
            for (unsigned e = 0; e < bSize; ++e) {
              const double x = want.dat.get()[b + e];
              worst = std::max(worst, std::fabs(got.dat.get()[b + e] - x) / std::max(std::fabs(x), 1e-300));
            }
          }
      std::cout << "  | max relative error vs naive kernel " << worst << std::endl;
      if (!(worst <= 1e-12)) {
        std::cout << "result mismatch!" << std::endl;
        ++failures;
      }
    }
    free(in_ptr);
    free(out_ptr);
    free(bInfo.adj);
    free(grid_ptr);
  });
  return rc ? rc : (failures ? 1 : 0);
}
