// Device-side building blocks of the marching stencil kernels (sm_100a): mbarrier + TMA
// (cp.async.bulk.tensor, SASS UTMALDG) wrappers, the per-stage shared-memory layout and the
// per-lane addressing constants.
//
// Work decomposition (see DESIGN.md "Kernels"): ONE WARP marches a column of brick layers
// along k for a footprint of two bricks adjacent in i (16 x 8 points per plane).  Lane
// (bb, j, ih) owns the four consecutive points i = 4*ih .. 4*ih+3 of row j of brick bb, so a
// plane of one brick is read by 16 lanes as 32-byte quads (two LDS.128), k-neighbours live in
// registers of the marching lane, i-neighbours come from the adjacent lane by warp shuffle
// (the `dev_shl` idea of the reference, include/dev_shl.h:51-63) and j-neighbours from shared
// memory.  Each stage of the 2-deep TMA pipeline holds P planes of: the two centre bricks, the
// i-face columns of the -i/+i neighbours, the j-face rows of the -j/+j neighbours and (cube
// stencil) the four ij-corner pieces; every piece is one TMA box copy out of the neighbour
// brick that BrickInfo::adj names (include/brick.h:96-127).
#pragma once
#include "common.cuh"

namespace dev {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
#ifdef MARCH_WAIT_TRAP
/* experiment builds (tools/build_variants.sh adds the flag): a stage that never arrives -- a
 * bookkeeping bug in a new pipeline -- ends in a trap after ~1 s instead of hanging the GPU */
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  const long long t0 = clock64();
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
    if (clock64() - t0 > 2000000000ll) __trap();
  }
}
#else
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
#endif
/* true in exactly one lane of the (converged) warp; the compiler knows it, so warp-uniform
 * instructions behind it (UTMALDG) need no loop over the active lanes */
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
/* one box of the [brick][k][j][i] field tensor -> dense box in shared memory */
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, int ci, int cj,
                                            int ck, int cb, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
      "l"(map), "r"(ci), "r"(cj), "r"(ck), "r"(cb), "r"(bar)
      : "memory");
}
/* 16 bytes global -> shared through the LSU (SASS LDGSTS), L2 only */
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void st_quad(double *p, const double (&v)[4]) {
  /* 32 contiguous bytes per lane = one full L2 sector, streaming (written once per sweep), as ONE
   * 256-bit store (SASS STG.E.EF.256): half the store instructions and L1TEX tag requests of two
   * 128-bit stores, and the sector reaches L2 whole instead of as two 16-byte partial writes
   * (measured: 125-point sweep 0.606 -> 0.581 ms, profiles/r02_cube_experiments.md) */
  asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3])
               : "memory");
}

/* ld.global.nc that the compiler may not sink to its use: adjacency rows are fetched a whole
 * stage before the copies that need them */
__device__ __forceinline__ unsigned ld_nc_pinned(const unsigned *p) {
  unsigned v;
  asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

/* predicated 256-bit streaming store without a branch and without a compiler memory barrier: the
 * planes of a stage stay one basic block, so ptxas can start the next plane's shared-memory loads
 * under the current plane's arithmetic (nothing in these kernels reads `out`) */
__device__ __forceinline__ void st_quad_if(double *p, const double (&v)[4], bool on) {
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "setp.ne.b32 q, %5, 0;\n"
      "@q st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};\n"
      "}\n" ::"l"(p),
      "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]), "r"((int)on));
}
/* fp32 bricks: a lane's quad is 16 bytes, one 128-bit streaming store; a warp row still covers whole lines */
__device__ __forceinline__ void st_quad_if(float *p, const float (&v)[4], bool on) {
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "setp.ne.b32 q, %5, 0;\n"
      "@q st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};\n"
      "}\n" ::"l"(p),
      "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "r"((int)on));
}
/* cp.async without the compiler memory barrier: ordering against the loads of a stage comes from
 * cp_async_wait / __syncwarp, which keep theirs */
__device__ __forceinline__ void cp_async16_nb(uint32_t dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
}

/* Byte layout of one pipeline stage.  P = planes per main stage (4 or 8); T = element type (double, or
 * float for fp32 bricks: 32-byte rows, 2 KiB bricks). */
template <int R, bool CUBE, int P, typename T = double>
struct StageLayout {
  static constexpr int EB = (int)sizeof(T);     /* element bytes */
  static constexpr int HW = (R > 2 ? 4 : 2) * EB >= 16 ? (R > 2 ? 4 : 2) : 16 / EB; /* i-face width in elements (16-byte multiple) */
  static constexpr int C_PL = 64 * EB, I_PL = 8 * EB * HW, J_PL = 8 * EB * R, X_PL = EB * R * HW; /* B/plane */
  /* 25-point star (R = 4): both j-face boxes are four rows copied with SWIZZLE_64B, which stores the
   * rows with their 16-byte chunks XORed by the row pair index (plus 2 in odd planes): face rows
   * -4..-1 and 8..11 then continue the bank pattern of the centre image, and the two-way bank
   * conflicts between face rows and centre rows are gone (ncu: 26 % of the kernel's shared-memory
   * wavefronts were conflicts, L1TEX 90 % busy; 0.613 -> 0.542 ms per 66^3-brick sweep,
   * profiles/r03_experiments/variants_k_prepared.log).  For R = 2 (13-point star, cube) the same
   * trick needs a four-row box for a two-row face and measured slower, so it is not used there. */
  static constexpr bool JLX = !CUBE && R == 4 && EB == 8; /* fp32 rows are 32 bytes: four rows per 128-byte line, no swizzle */
  static constexpr bool JHX = JLX;
  static constexpr int JL_ROWS = JLX ? 4 : R, JL_PL = 8 * EB * JL_ROWS;
  static constexpr int JSUM_PL = 2 * J_PL + 2 * JL_PL; /* the four j-face pieces, bytes per plane */
  static constexpr int C0 = 0, C1 = P * C_PL;
  static constexpr int IL = 2 * P * C_PL, IR = IL + P * I_PL;
  static constexpr int JL0 = IR + P * I_PL, JL1 = JL0 + P * JL_PL, JH0 = JL1 + P * JL_PL,
                       JH1 = JH0 + P * J_PL;
  static_assert(!JHX || (JH0 % 512 == 0 && (P * J_PL) % 512 == 0), "swizzled +j face pieces: 512-byte phase");
  static_assert(!JLX || (JL0 % 512 == 0 && (P * JL_PL) % 512 == 0 && (R * JL_PL) % 512 == 0),
                "swizzled -j face pieces must keep the 512-byte phase of the swizzle pattern");
  /* a TMA destination must be 128-byte aligned: the 2x1 corner pieces of the 27-point cube are
   * only 16 bytes per plane */
  static constexpr int X_STRIDE = P * X_PL < 128 ? 128 : P * X_PL;
  static constexpr int XLL = JH1 + P * J_PL, XRL = XLL + X_STRIDE, XLH = XRL + X_STRIDE,
                       XRH = XLH + X_STRIDE;
  static constexpr int RAW = CUBE ? XRH + X_STRIDE : XLL;
  static constexpr int BYTES = (RAW + 511) / 512 * 512;
  static constexpr int NPIECE_MAIN = CUBE ? 12 : 8;
  static constexpr int NPIECE_ROLL = CUBE ? 12 : 2;
};

/* what lane p (< 12) copies: piece p of the stage */
struct PieceK {
  int useb1, di, dj, map, ci, cj, off, plane_bytes, pairmap;
};

template <int R, bool CUBE, int P, typename T = double>
__device__ __forceinline__ PieceK piece_of(int p) {
  using L = StageLayout<R, CUBE, P, T>;
  PieceK k;
  k.useb1 = p & 1;
  const int grp = p >> 1; /* 0 C, 1 I, 2 JL, 3 JH, 4 X-low (LL,RL), 5 X-high (LH,RH) */
  const int side = k.useb1 ? +1 : -1;
  k.di = (grp == 1 || grp >= 4) ? side : 0;
  k.dj = (grp == 2 || grp == 4) ? -1 : ((grp == 3 || grp == 5) ? +1 : 0);
  k.map = grp == 0 ? 0 : (grp == 1 ? 1 : (grp <= 3 ? 2 : 3));
  k.pairmap = grp == 0 ? 0 : 1;
  k.ci = (k.di < 0) ? 8 - L::HW : 0;
  k.cj = (k.dj < 0) ? 8 - R : 0;
  k.plane_bytes = grp == 0 ? L::C_PL : (grp == 1 ? L::I_PL : (grp <= 3 ? L::J_PL : L::X_PL));
  if (L::JHX && grp == 3) { /* same box shape at rows 0..3 of the +j neighbour */
    k.map = 4;
    k.pairmap = 2;
  }
  if (L::JLX && grp == 2) { /* the swizzled four-row -j face box has its own maps */
    k.map = 4;
    k.pairmap = 2;
    k.cj = 8 - L::JL_ROWS;
    k.plane_bytes = L::JL_PL;
  }
  const int offs[12] = {L::C0,  L::C1,  L::IL,  L::IR,  L::JL0, L::JL1,
                        L::JH0, L::JH1, L::XLL, L::XRL, L::XLH, L::XRH};
  k.off = offs[p < 12 ? p : 0];
  return k;
}

/* Per-lane addressing constants.  Row j+d of the lane's own brick lives either in the centre
 * piece (64B-swizzled by TMA: 16-byte chunk index ^= (row >> 1) & 3) or in a j-face piece. */
template <int R, bool CUBE, int P, typename T = double>
struct LaneK {
  using L = StageLayout<R, CUBE, P, T>;
  int bb, j, ih;
  int srcL, srcR;        /* shuffle sources for the left / right i-neighbour quads */
  bool halo;             /* this lane loads the i-face (left if bb==0, right if bb==1) */
  int row0[2 * R + 1];   /* byte offset of quad chunk 0 of row j+d, plane 0 */
  int row1[2 * R + 1];   /* ... chunk 1 */
  int rpitch[2 * R + 1]; /* bytes per plane for that row */
  int hrow[2 * R + 1];   /* halo: byte offset of the HW-wide row j+d, plane 0 */
  int hpitch[2 * R + 1];
  int jx[2 * R + 1]; /* XOR for the quad offsets of row j+d in ODD planes: 0x20 when it is a swizzled face row */

  __device__ __forceinline__ void init(int lane) {
    bb = lane >> 4;
    j = (lane >> 1) & 7;
    ih = lane & 1;
    halo = (ih == bb);
    srcL = ih ? lane - 1 : (bb ? lane - 15 : lane);
    srcR = ih ? (bb ? lane : lane + 15) : lane + 1;
#pragma unroll
    for (int d = -R; d <= R; ++d) {
      const int jj = j + d, a = d + R;
      jx[a] = 0;
      if constexpr (sizeof(T) == 4) {
        /* fp32: rows of 32 bytes, the lane's quad is the 16-byte half `ih` of its row; the eight lanes of
         * a quarter-warp read four consecutive rows = 128 contiguous bytes, so nothing is swizzled */
        constexpr int ROWB = 8 * L::EB, HROWB = L::HW * L::EB;
        if (jj >= 0 && jj < 8) {
          row0[a] = (bb ? L::C1 : L::C0) + jj * ROWB + ih * 16;
          rpitch[a] = L::C_PL;
          hrow[a] = (bb ? L::IR : L::IL) + jj * HROWB;
          hpitch[a] = L::I_PL;
        } else {
          const bool low = jj < 0;
          const int rr = low ? jj + R : jj - 8;
          row0[a] = (low ? (bb ? L::JL1 : L::JL0) : (bb ? L::JH1 : L::JH0)) + rr * ROWB + ih * 16;
          rpitch[a] = L::J_PL;
          hrow[a] = (low ? (bb ? L::XRL : L::XLL) : (bb ? L::XRH : L::XLH)) + rr * HROWB;
          hpitch[a] = L::X_PL;
        }
        row1[a] = row0[a];
        continue;
      }
      if (L::JHX && jj > 7) {
        const int rr = jj - 8;
        const int base = (bb ? L::JH1 : L::JH0) + rr * 64, x0 = (rr >> 1) & 3;
        row0[a] = base + (((2 * ih) ^ x0) << 4);
        row1[a] = base + (((2 * ih + 1) ^ x0) << 4);
        rpitch[a] = L::J_PL;
        jx[a] = 0x20;
        hrow[a] = (bb ? L::XRH : L::XLH) + rr * L::HW * 8;
        hpitch[a] = L::X_PL;
        continue;
      }
      if (L::JLX && jj < 0) {
        /* row jj of the brick = row rr of the four-row box; SWIZZLE_64B XORs the 16-byte chunk index
         * with bits 7..8 of the shared-memory address: (rr >> 1) + 2 * plane for a 512-aligned piece */
        const int rr = jj + L::JL_ROWS;
        const int base = (bb ? L::JL1 : L::JL0) + rr * 64, x0 = (rr >> 1) & 3;
        row0[a] = base + (((2 * ih) ^ x0) << 4);
        row1[a] = base + (((2 * ih + 1) ^ x0) << 4);
        rpitch[a] = L::JL_PL;
        jx[a] = 0x20; /* odd planes: chunk ^ (x0 ^ 2) */
        hrow[a] = (bb ? L::XRL : L::XLL) + (jj + R) * L::HW * 8;
        hpitch[a] = L::X_PL;
        continue;
      }
      if (jj >= 0 && jj < 8) {
        const int base = (bb ? L::C1 : L::C0) + jj * 64, sw = (jj >> 1) & 3;
        row0[a] = base + (((2 * ih) ^ sw) << 4);
        row1[a] = base + (((2 * ih + 1) ^ sw) << 4);
        rpitch[a] = L::C_PL;
        hrow[a] = (bb ? L::IR : L::IL) + jj * L::HW * 8;
        hpitch[a] = L::I_PL;
      } else {
        const bool low = jj < 0;
        const int rr = low ? jj + R : jj - 8;
        const int base = (low ? (bb ? L::JL1 : L::JL0) : (bb ? L::JH1 : L::JH0)) + rr * 64;
        row0[a] = base + ih * 32;
        row1[a] = base + ih * 32 + 16;
        rpitch[a] = L::J_PL;
        hrow[a] = (low ? (bb ? L::XRL : L::XLL) : (bb ? L::XRH : L::XLH)) + rr * L::HW * 8;
        hpitch[a] = L::X_PL;
      }
    }
  }
};

__device__ __forceinline__ void ld_quad(const char *st, int off0, int /*off1*/, float (&v)[4]) {
  const float4 a = *reinterpret_cast<const float4 *>(st + off0); /* the whole quad is one 16-byte chunk */
  v[0] = a.x;
  v[1] = a.y;
  v[2] = a.z;
  v[3] = a.w;
}
__device__ __forceinline__ void ld_quad(const char *st, int off0, int off1, double (&v)[4]) {
  const double2 a = *reinterpret_cast<const double2 *>(st + off0);
  const double2 b = *reinterpret_cast<const double2 *>(st + off1);
  v[0] = a.x;
  v[1] = a.y;
  v[2] = b.x;
  v[3] = b.y;
}

}  // namespace dev
