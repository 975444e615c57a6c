/*
 * oracle/decomp_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's brick numbering for 3-D, 8x8x8 bricks:
 *   - BrickDecomp::initialize        include/brick-mpi.h:322-460
 *     (_populate :218-239, populate :241-254, get_region_size :256-264,
 *      _adj_populate/adj_populate :266-291)
 *   - allneighbors                   src/brick-mpi.cpp:9-23
 *   - populate() rank map            include/brick-mpi.h:730-753 (+ MPI-standard row-major,
 *                                    periodic MPI_Cart_rank)
 *   - init_grid / init_fill          include/bricksetup.h:30-90
 * PINNED: tests/test_oracle_golden.py compares every table it produces with
 * tests/golden/decomp_*.npz (dumped from the compiled reference by oracle/gen_golden.py) and,
 * when oracle/_ref is present, with the reference itself at 512^3.
 * Set encoding (include/bitset.h:18-40): +d -> bit d, -d -> bit 31+d.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  uint64_t neighbor;
  unsigned skin_st, skin_ed, pos, len, first_pad, last_pad;
} oracle_region;

typedef struct {
  unsigned dims[3], tdims[3], gdepth[3], stride[3];
  unsigned nbricks, sep_pos[3];
  unsigned *grid; /* tdims[2]*tdims[1]*tdims[0], i fastest */
  unsigned *adj;  /* nbricks * 27 */
  int nregions;
  oracle_region *ghost, *skin;
  int nskin;
  long *skin_size;
} oracle_decomp;

static uint64_t bit_of(int d) { return 1ull << (d < 0 ? 31 - d : d); }
static int has(uint64_t s, int d) { return (s & bit_of(d)) != 0; }
static int popcnt(uint64_t s) { return __builtin_popcountll(s); }
static uint64_t negate(uint64_t s) {
  uint64_t lo = s & ((1ull << 32) - 1);
  return (lo << 31) | (s >> 31);
}

static void number_region(oracle_decomp *D, uint64_t region, long ref, int d, unsigned *pos) {
  if (d < 0) {
    D->grid[ref] = (*pos)++;
    return;
  }
  long lo, n;
  if (has(region, -(d + 1))) { /* the reference tests + first, then - overrides */
    lo = D->gdepth[d];
    n = D->gdepth[d];
  } else if (has(region, d + 1)) {
    lo = D->dims[d];
    n = D->gdepth[d];
  } else {
    lo = 2 * D->gdepth[d];
    n = (long)D->dims[d] - lo;
  }
  for (long i = 0; i < n; ++i) number_region(D, region, ref + (lo + i) * D->stride[d], d - 1, pos);
}

static void number_owned(oracle_decomp *D, uint64_t owner, uint64_t region, unsigned *pos) {
  long ref = 0;
  for (int d = 0; d < 3; ++d) {
    if (has(owner, d + 1)) ref += (long)D->dims[d] * D->stride[d];
    if (has(owner, -(d + 1))) ref -= (long)D->dims[d] * D->stride[d];
  }
  number_region(D, region, ref, 2, pos);
}

static long region_size(const oracle_decomp *D, uint64_t region) {
  long r = 1;
  for (int d = 1; d <= 3; ++d)
    r *= (has(region, d) || has(region, -d)) ? D->gdepth[d - 1] : D->dims[d - 1] - 2 * D->gdepth[d - 1];
  return r;
}

int oracle_allneighbors(uint64_t *out) {
  int n = 0;
  static const int sgn[3] = {+1, 0, -1};
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b)
      for (int c = 0; c < 3; ++c) {
        uint64_t s = 0;
        if (sgn[a]) s |= bit_of(sgn[a] * 1);
        if (sgn[b]) s |= bit_of(sgn[b] * 2);
        if (sgn[c]) s |= bit_of(sgn[c] * 3);
        out[n++] = s;
      }
  return n;
}

static void fill_adj(const oracle_decomp *D, long cell, unsigned *row) {
  long c[3] = {cell % D->tdims[0], (cell / D->stride[1]) % D->tdims[1], (cell / D->stride[2]) % D->tdims[2]};
  for (int ok = 0; ok < 3; ++ok)
    for (int oj = 0; oj < 3; ++oj)
      for (int oi = 0; oi < 3; ++oi) {
        long k = c[2] + ok - 1, j = c[1] + oj - 1, i = c[0] + oi - 1;
        int inside = k >= 0 && k < D->tdims[2] && j >= 0 && j < D->tdims[1] && i >= 0 && i < D->tdims[0];
        row[(ok * 3 + oj) * 3 + oi] = inside ? D->grid[k * D->stride[2] + j * D->stride[1] + i] : 0;
      }
}

/* dims: subdomain extents in ELEMENTS, contiguous first; depth: ghost depth in elements;
 * skinlist: nskin sets; page_align = 0 reproduces -DDECOMP_PAGEUNALIGN (weak/CMakeLists.txt:32);
 * pagesize/brick_bytes give `factor` (brick-mpi.h:342-350). */
oracle_decomp *oracle_decomp_build(const unsigned *dims, unsigned depth, const uint64_t *skinlist,
                                   int nskin, int page_align, long pagesize, long brick_bytes) {
  oracle_decomp *D = (oracle_decomp *)calloc(1, sizeof(oracle_decomp));
  unsigned cells = 1;
  for (int d = 0; d < 3; ++d) {
    D->dims[d] = dims[d] / 8;
    D->gdepth[d] = depth / 8;
    D->stride[d] = cells;
    D->tdims[d] = D->dims[d] + 2 * D->gdepth[d];
    cells *= D->tdims[d];
  }
  D->grid = (unsigned *)calloc(cells, sizeof(unsigned));
  long factor = brick_bytes < pagesize ? pagesize / brick_bytes : 1;
  unsigned pos = (unsigned)factor;
#define ALIGN_POS() do { if (page_align) pos = (unsigned)((pos + factor - 1) / factor * factor); } while (0)
#define PAD_OF(reg) (page_align ? (factor - (region_size(D, (reg)) + factor - 1) % factor - 1) : 0)

  int *pad_first = (int *)calloc(nskin, sizeof(int));
  unsigned *st_pos = (unsigned *)calloc(nskin + 1, sizeof(unsigned));
  uint64_t last = 0;
  for (int i = 0; i < nskin; ++i) {
    uint64_t next = i + 1 < nskin ? skinlist[i + 1] : 0;
    pad_first[i] = popcnt(last & skinlist[i]) < popcnt(skinlist[i] & next);
    last = skinlist[i];
  }
  number_owned(D, 0, 0, &pos);
  ALIGN_POS();
  st_pos[0] = pos;
  D->sep_pos[0] = pos;
  D->nskin = nskin;
  D->skin_size = (long *)calloc(nskin, sizeof(long));
  for (int i = 0; i < nskin; ++i) {
    unsigned before = pos;
    if (pad_first[i]) pos += (unsigned)PAD_OF(skinlist[i]);
    if (skinlist[i]) {
      number_owned(D, 0, skinlist[i], &pos);
      ALIGN_POS();
    }
    st_pos[i + 1] = pos;
    D->skin_size[i] = (long)pos - before;
  }
  D->sep_pos[1] = pos;

  uint64_t nbs[27];
  oracle_allneighbors(nbs);
  D->ghost = (oracle_region *)calloc(27 * (nskin + 1), sizeof(oracle_region));
  D->skin = (oracle_region *)calloc(27 * (nskin + 1), sizeof(oracle_region));
  int nr = 0;
  for (int q = 0; q < 27; ++q) {
    uint64_t n = nbs[q];
    if (!n) continue;
    uint64_t in = negate(n);
    oracle_region g, s;
    memset(&g, 0, sizeof g);
    memset(&s, 0, sizeof s);
    g.neighbor = n;
    s.neighbor = in;
    int open = -1;
    for (int l = 0; l <= nskin; ++l) {
      int member = l < nskin && (in & skinlist[l]) == in;
      if (member) {
        if (open < 0) {
          open = l;
          g.skin_st = s.skin_st = (unsigned)l;
          g.first_pad = s.first_pad = pad_first[l] ? (unsigned)PAD_OF(skinlist[l]) : 0;
          g.pos = pos;
          s.pos = st_pos[l];
        }
        if (pad_first[l]) pos += (unsigned)PAD_OF(skinlist[l]);
        number_owned(D, n, skinlist[l], &pos);
        ALIGN_POS();
      } else if (open >= 0) {
        g.last_pad = s.last_pad = pad_first[l - 1] ? 0 : (unsigned)PAD_OF(skinlist[l - 1]);
        g.skin_ed = s.skin_ed = (unsigned)l;
        g.len = pos - g.pos;
        s.len = st_pos[l] - s.pos;
        D->ghost[nr] = g;
        D->skin[nr] = s;
        ++nr;
        open = -1;
      }
    }
  }
  D->nregions = nr;
  D->sep_pos[2] = pos;
  D->nbricks = pos;
  D->adj = (unsigned *)calloc((size_t)pos * 27, sizeof(unsigned));
  for (long c = 0; c < cells; ++c) fill_adj(D, c, D->adj + (size_t)D->grid[c] * 27);
  free(pad_first);
  free(st_pos);
  return D;
}

void oracle_decomp_free(oracle_decomp *D) {
  if (!D) return;
  free(D->grid);
  free(D->adj);
  free(D->ghost);
  free(D->skin);
  free(D->skin_size);
  free(D);
}

/* flat getters for ctypes */
unsigned oracle_decomp_nbricks(const oracle_decomp *D) { return D->nbricks; }
int oracle_decomp_nregions(const oracle_decomp *D) { return D->nregions; }
void oracle_decomp_get(const oracle_decomp *D, unsigned *sep_pos, unsigned *grid, unsigned *adj,
                       oracle_region *ghost, oracle_region *skin, long *skin_size) {
  size_t cells = (size_t)D->tdims[0] * D->tdims[1] * D->tdims[2];
  memcpy(sep_pos, D->sep_pos, sizeof D->sep_pos);
  if (grid) memcpy(grid, D->grid, cells * sizeof(unsigned));
  if (adj) memcpy(adj, D->adj, (size_t)D->nbricks * 27 * sizeof(unsigned));
  memcpy(ghost, D->ghost, D->nregions * sizeof(oracle_region));
  memcpy(skin, D->skin, D->nregions * sizeof(oracle_region));
  memcpy(skin_size, D->skin_size, D->nskin * sizeof(long));
}

/* rank_map of populate(): keys in allneighbors order; procdims/coo as MPI_Cart (index 0 slowest) */
void oracle_rank_map(const int *procdims, const int *coo, uint64_t *keys, int *ranks) {
  static const int sgn[3] = {+1, 0, -1};
  int n = 0;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b)
      for (int c = 0; c < 3; ++c) {
        int s[3] = {sgn[a], sgn[b], sgn[c]}; /* axis 1 (i), 2 (j), 3 (k) */
        uint64_t key = 0;
        int cc[3] = {coo[0], coo[1], coo[2]};
        for (int d = 1; d <= 3; ++d) {
          if (s[d - 1]) key |= bit_of(s[d - 1] * d);
          cc[3 - d] = coo[3 - d] - s[d - 1]; /* +d -> coordinate - 1, -d -> + 1 */
        }
        int r = 0;
        for (int d = 0; d < 3; ++d) r = r * procdims[d] + ((cc[d] % procdims[d]) + procdims[d]) % procdims[d];
        keys[n] = key;
        ranks[n] = r;
        ++n;
      }
}

/* init_grid<3>: identity numbering, out-of-range neighbours -> 0 (bricksetup.h:30-90) */
void oracle_init_grid(const long *dimlist, unsigned *grid, unsigned *adj) {
  long n0 = dimlist[0], n1 = dimlist[1], n2 = dimlist[2];
  long cells = n0 * n1 * n2;
  for (long p = 0; p < cells; ++p) grid[p] = (unsigned)p;
  for (long p = 0; p < cells; ++p)
    for (int ok = -1; ok <= 1; ++ok)
      for (int oj = -1; oj <= 1; ++oj)
        for (int oi = -1; oi <= 1; ++oi) {
          /* the reference bounds-checks the flat pointer only (init_fill, StopTag) */
          long q = p + oi + oj * n0 + ok * n0 * n1;
          adj[p * 27 + ((ok + 1) * 3 + oj + 1) * 3 + oi + 1] = (q >= 0 && q < cells) ? (unsigned)q : 0;
        }
}
