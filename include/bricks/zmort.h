// zmort.h -- n-D Z-order (Morton) index.
//
// Drop-in for the reference's include/zmort.h:18-105: `ZMORT()[a][b][c]` builds the index one
// coordinate at a time (the coordinate appended LAST occupies the least significant bit of
// every bit group), `z(d)` extracts the coordinate held at bit `d` of each group, `z.set(d,p)`
// replaces it.  Used by the strong-scaling driver to name subdomains (strong/args.cpp:36-56).
#ifndef BRICKS_B200_ZMORT_H
#define BRICKS_B200_ZMORT_H

struct ZMORT {
  unsigned long id;   ///< the interleaved index
  unsigned long dim;  ///< number of coordinates interleaved so far

  ZMORT() : id(0ul), dim(0ul) {}
  ZMORT(unsigned long id, unsigned long dim) : id(id), dim(dim) {}

  /// append coordinate p as the new least significant member of each group
  inline ZMORT operator[](unsigned long p) const {
    ZMORT out(0ul, dim + 1ul);
    const unsigned long group = dim ? ((1ul << dim) - 1ul) : 0ul;
    unsigned long rest = id;
    for (unsigned long level = 0; rest || p; ++level) {
      const unsigned long bits = ((rest & group) << 1ul) | (p & 1ul);
      out.id |= bits << (level * out.dim);
      rest = dim ? rest >> dim : 0ul;
      p >>= 1ul;
    }
    return out;
  }

  /// coordinate stored at bit d of each group
  inline unsigned long operator()(unsigned long d) const {
    unsigned long coord = 0ul;
    unsigned long level = 0ul;
    for (unsigned long rest = id >> d; rest; rest >>= dim, ++level) coord |= (rest & 1ul) << level;
    return coord;
  }

  /// copy with coordinate d replaced by p
  inline ZMORT set(unsigned long d, unsigned long p) const {
    ZMORT out(0ul, dim);
    const unsigned long keep = ((1ul << dim) - 1ul) & ~(1ul << d);  // the other coordinates' bits
    unsigned long rest = id;
    for (unsigned long shift = 0; rest || p; shift += dim) {
      out.id |= ((rest & keep) | ((p & 1ul) << d)) << shift;
      rest >>= dim;
      p >>= 1ul;
    }
    return out;
  }

  inline operator unsigned long() const { return id; }
};

#endif  // BRICKS_B200_ZMORT_H
