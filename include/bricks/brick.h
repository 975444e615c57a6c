// brick.h -- the brick data layout: storage, adjacency metadata and the indexable Brick view.
//
// Drop-in for the reference's include/brick.h (BrickStorage :53-82, BrickInfo :96-127,
// Dim/cal_size/cal_offs :130-194, _BrickAccessor :214-327, Brick :353-395) and the bElem
// typedef of include/vecscatter.h:12-14.  Same type names, member names and memory layout,
// so code written against bricklib compiles against this header; the implementation is
// new (one accessor template driven by constexpr tables instead of three partial
// specialisations).  Host only: device code never sees these types, it gets raw
// adj/dat/step through the C ABI (include/bricks_b200.h).
//
// Layout contract (the thing the CUDA kernels rely on): element (b; x_{n-1} .. x_0) of a
// field lives at dat[adj-resolved-brick * step + vec * VECLEN + lane] where, walking the
// axes from the outermost to x_0 with fold F per axis (1 where no fold is given),
//   vec  = vec  * (D / F) + x / F        lane = lane * F + x % F.
// With fold <4,8> or <8> on 8x8x8 bricks this is plain row-major k*64 + j*8 + i.
#ifndef BRICKS_B200_BRICK_H
#define BRICKS_B200_BRICK_H

#include <cstddef>
#include <cstdlib>
#include <memory>
#include <type_traits>

#ifndef bElem
#define bElem double
#endif

/// BrickStorage allocation alignment (bytes)
#ifndef ALIGN
#define ALIGN 2048
#endif

#ifndef FORCUDA
#define FORCUDA
#endif

template <unsigned base, unsigned exp>
struct static_power {
  static constexpr unsigned value = base * static_power<base, exp - 1>::value;
};
template <unsigned base>
struct static_power<base, 0> {
  static constexpr unsigned value = 1;
};

/// Owns (or shares) the memory of `chunks` chunks of `step` elements each.  A chunk holds one
/// brick of every interleaved field.
struct BrickStorage {
  std::shared_ptr<bElem> dat;
  long chunks = 0;
  size_t step = 0;
  void *mmap_info = nullptr;  ///< kept for source compatibility; this build never mmaps

  static BrickStorage allocate(long chunks, size_t step) {
    BrickStorage s;
    s.chunks = chunks;
    s.step = step;
    size_t bytes = (size_t)chunks * step * sizeof(bElem);
    bytes = (bytes + ALIGN - 1) / ALIGN * ALIGN;  // aligned_alloc wants a multiple
    s.dat = std::shared_ptr<bElem>((bElem *)aligned_alloc(ALIGN, bytes), [](bElem *p) { free(p); });
    return s;
  }
};

/// Adjacency metadata: adj[b][n] is the storage index of neighbour n of brick b, n written in
/// base 3 with one digit per axis (outermost axis most significant, digit 1 = same brick).
template <unsigned dims>
struct BrickInfo {
  typedef unsigned (*adjlist)[static_power<3, dims>::value];
  adjlist adj;
  unsigned nbricks;

  explicit BrickInfo(unsigned nbricks) : nbricks(nbricks) {
    adj = (adjlist)malloc((size_t)nbricks * static_power<3, dims>::value * sizeof(unsigned));
  }
  BrickStorage allocate(long step) { return BrickStorage::allocate(nbricks, step); }
};

template <unsigned... Ds>
struct Dim {};

template <unsigned... xs>
struct cal_size;
template <unsigned x>
struct cal_size<x> {
  static constexpr unsigned value = x;
};
template <unsigned x, unsigned... xs>
struct cal_size<x, xs...> {
  static constexpr unsigned value = x * cal_size<xs...>::value;
};

/// cal_offs<n, o_{n-1}, .., o_0>: adjacency slot of the neighbour with base-3 digits o
template <unsigned... offs>
struct cal_offs;
template <unsigned off>
struct cal_offs<1, off> {
  static constexpr unsigned value = off;
};
template <unsigned dim, unsigned off, unsigned... offs>
struct cal_offs<dim, off, offs...> {
  static constexpr unsigned value = off * static_power<3, dim - 1>::value + cal_offs<dim - 1, offs...>::value;
};

namespace bricks_detail {
template <unsigned N>
struct UTable {
  unsigned v[N];
};
template <unsigned N, unsigned... Xs>
constexpr UTable<N> right_aligned(unsigned fill) {
  UTable<N> t{};
  for (unsigned i = 0; i < N; ++i) t.v[i] = fill;
  const unsigned src[sizeof...(Xs) ? sizeof...(Xs) : 1] = {Xs...};
  for (unsigned i = 0; i < sizeof...(Xs); ++i) t.v[N - sizeof...(Xs) + i] = src[i];
  return t;
}
}  // namespace bricks_detail

template <typename...>
struct Brick;

/// One level of `brick[b][k][j][i]`: AXIS axes have been consumed so far.
template <typename B, unsigned AXIS>
struct BrickAccessorLevel {
  B *par;
  unsigned b, pos, nvec, wvec;

  typedef typename std::conditional<AXIS + 1 == B::NDIMS, bElem &, BrickAccessorLevel<B, AXIS + 1>>::type next_t;

  FORCUDA inline next_t operator[](unsigned x) { return step(x, std::integral_constant<bool, AXIS + 1 == B::NDIMS>()); }

 private:
  // x may be "negative" (unsigned wrap) or >= D: x + D lands in [0, 3D) and its digit picks
  // the neighbour (include/brick.h:229-246 of the reference defines exactly this arithmetic)
  FORCUDA inline void advance(unsigned x, unsigned &npos, unsigned &n, unsigned &w) const {
    constexpr unsigned D = B::bdims.v[AXIS], F = B::folds.v[AXIS];
    const unsigned dir = x + D;
    npos = pos * 3 + dir / D;
    const unsigned l = dir % D;
    n = nvec * (D / F) + l / F;
    w = wvec * F + l % F;
  }
  FORCUDA inline bElem &step(unsigned x, std::true_type) {
    unsigned npos, n, w;
    advance(x, npos, n, w);
    return par->dat[(size_t)par->bInfo->adj[b][npos] * par->step + n * B::VECLEN + w];
  }
  FORCUDA inline BrickAccessorLevel<B, AXIS + 1> step(unsigned x, std::false_type) {
    unsigned npos, n, w;
    advance(x, npos, n, w);
    return BrickAccessorLevel<B, AXIS + 1>{par, b, npos, n, w};
  }
};

template <unsigned... BDims, unsigned... Folds>
struct Brick<Dim<BDims...>, Dim<Folds...>> {
  typedef Brick<Dim<BDims...>, Dim<Folds...>> mytype;
  typedef BrickInfo<sizeof...(BDims)> myBrickInfo;

  static constexpr unsigned NDIMS = sizeof...(BDims);
  static constexpr unsigned VECLEN = cal_size<Folds...>::value;
  static constexpr unsigned BRICKSIZE = cal_size<BDims...>::value;
  static constexpr bricks_detail::UTable<NDIMS> bdims = bricks_detail::right_aligned<NDIMS, BDims...>(1);
  static constexpr bricks_detail::UTable<NDIMS> folds = bricks_detail::right_aligned<NDIMS, Folds...>(1);
  static_assert(sizeof...(Folds) <= sizeof...(BDims), "more fold axes than brick axes");

  myBrickInfo *bInfo;     ///< (possibly shared) adjacency metadata
  size_t step;            ///< elements between consecutive bricks of this field
  bElem *dat;             ///< storage base + field offset
  BrickStorage bStorage;  ///< keeps the memory alive

  Brick(myBrickInfo *bInfo, const BrickStorage &storage, unsigned offset)
      : bInfo(bInfo), step((unsigned)storage.step), dat(storage.dat.get() + offset), bStorage(storage) {}

  FORCUDA inline BrickAccessorLevel<mytype, 0> operator[](unsigned b) {
    return BrickAccessorLevel<mytype, 0>{this, b, 0, 0, 0};
  }

  /// first element of the neighbour brick whose base-3 digits are Offsets (outermost first)
  template <unsigned... Offsets>
  FORCUDA inline bElem *neighbor(unsigned b) {
    return &dat[(size_t)bInfo->adj[b][cal_offs<sizeof...(BDims), Offsets...>::value] * step];
  }

  /// true when the in-brick order is plain row-major (the layout the CUDA kernels require)
  static constexpr bool row_major() {
    // every fold axis must cover its whole brick axis, except that the outermost folded axis
    // may be any divisor: then vec/lane arithmetic collapses to row-major
    bool seen_partial = false;
    for (unsigned a = NDIMS; a-- > 0;) {
      if (folds.v[a] == bdims.v[a]) {
        if (seen_partial) return false;
      } else if (folds.v[a] == 1) {
        seen_partial = true;
      } else {
        if (seen_partial) return false;
        seen_partial = true;
      }
    }
    return true;
  }
};

#if __cplusplus < 201703L
#error "include/bricks/*.h need C++17 (inline variables, constexpr tables)"
#endif

#endif  // BRICKS_B200_BRICK_H
