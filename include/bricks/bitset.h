// bitset.h -- set of signed axis ids naming a region / neighbour direction.
//
// Drop-in for the reference's include/bitset.h:18-177 (same type name, same public members,
// same bit encoding so `.set` values and rank_map keys are interchangeable):
//   axis +d (d >= 1)  -> bit d          axis -d -> bit 31 + d
// Axis 1 is i (contiguous), 2 is j, 3 is k (include/brick-mpi.h:218-239).
#ifndef BRICKS_B200_BITSET_H
#define BRICKS_B200_BITSET_H

#include <cctype>
#include <cstdint>
#include <initializer_list>
#include <istream>
#include <ostream>
#include <stdexcept>
#include <string>

struct BitSet {
  static const long zero = 31;  // where the negative axes start
  uint64_t set;

  static inline uint64_t to_set(long axis) {
    const long bit = axis >= 0 ? axis : zero - axis;
    return uint64_t(1) << uint64_t(bit);
  }

  BitSet() : set(0) {}
  BitSet(uint64_t bits) : set(bits) {}
  BitSet(std::initializer_list<int> axes) : set(0) {
    for (int a : axes) set ^= to_set(a);
  }

  inline BitSet &flip(long axis) {
    set ^= to_set(axis);
    return *this;
  }
  inline bool get(long axis) const { return (set & to_set(axis)) != 0; }
  inline long size() const { return __builtin_popcountll(set); }

  inline BitSet operator&(BitSet o) const { return BitSet(set & o.set); }
  inline BitSet operator|(BitSet o) const { return BitSet(set | o.set); }
  inline BitSet operator^(BitSet o) const { return BitSet(set ^ o.set); }
  inline operator bool() const { return set != 0; }
  /// subset / superset
  inline bool operator<=(BitSet o) const { return (set & ~o.set) == 0; }
  inline bool operator>=(BitSet o) const { return (o.set & ~set) == 0; }
  /// mirror every direction: +d <-> -d
  inline BitSet operator!() const {
    const uint64_t low = set & ((uint64_t(1) << (zero + 1)) - 1);
    return BitSet((low << zero) | (set >> zero));
  }
};

// "{1+2-}" form, identical to the reference's stream format (bitset.h:139-177)
inline std::ostream &operator<<(std::ostream &os, const BitSet &bs) {
  os << "{";
  for (long d = 1; d < 32; ++d) {
    if (bs.get(d)) os << d << "+";
    if (bs.get(-d)) os << d << "-";
  }
  return os << "}";
}

inline std::istream &operator>>(std::istream &is, BitSet &bs) {
  std::string tok;
  is >> tok;
  if (tok.empty() || tok[0] != '{') throw std::runtime_error("Err");
  BitSet out;
  size_t p = 1;
  while (p < tok.size() && tok[p] != '}') {
    long d = 0;
    while (p < tok.size() && isdigit((unsigned char)tok[p])) d = d * 10 + (tok[p++] - '0');
    if (p >= tok.size() || (tok[p] != '+' && tok[p] != '-')) throw std::runtime_error("Err");
    out.flip(tok[p] == '-' ? -d : d);
    ++p;
  }
  bs = out;
  return is;
}

#endif  // BRICKS_B200_BITSET_H
