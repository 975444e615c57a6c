"""oracle/oracle4d.py -- TEST INFRASTRUCTURE, not product code.

The 4-D path of the reference (weak/main-4d.cpp: 4x4x4x4 bricks, ghost depth 8, the 9-point stencil of
stencils/mpi9pt.py), restated for the checker:

  * DecompND: BrickDecomp<dim, B..>::initialize (include/brick-mpi.h:322-460 with _populate :218-239,
    populate :241-254, get_region_size :256-264, adj_populate :266-291) for any number of axes, in
    numpy -- integer work at test sizes.  Both numberings: page-aligned (the default build) and dense
    (-DDECOMP_PAGEUNALIGN).  allneighbors: src/brick-mpi.cpp:9-23; populate() rank map:
    include/brick-mpi.h:730-753.
  * brick_sweep / array_sweep / sweep_list / copy_brick: ctypes front end of oracle/stencil4d_oracle.c.
  * Ref4D / ref_brick_sweep: the compiled, unmodified reference (oracle/_ref/libref4d*.so), present in
    the build container and on the GPU box (the .so files travel), used to pin the above.

PINNED: tests/test_oracle_golden.py compares DecompND and brick_sweep with tests/golden/decomp4d.json
and tests/golden/d4pt9_*.npz (dumped from the compiled reference by `python oracle/gen_golden.py d4`),
and with the reference itself when oracle/_ref is present.
Set encoding (include/bitset.h:18-40): +d -> bit d, -d -> bit 31+d.
"""
import ctypes as C
import os

import numpy as np

from . import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
BDIM = 4
ELEMS = 256
NADJ = 81
# adjacency slots of the eight face neighbours: ((dl*3 + dk)*3 + dj)*3 + di, digit 1 = same brick
CENTRE = 40
FACE_SLOTS = {"i-": 39, "i+": 41, "j-": 37, "j+": 43, "k-": 31, "k+": 49, "l-": 13, "l+": 67}
# ghost depth in elements and sweeps per exchange of weak/main-4d.cpp:27-29,60
GZ = 8
ST_ITER = 8


def _bit(d):
    return 1 << (31 - d if d < 0 else d)


def has(s, d):
    return (s & _bit(d)) != 0


def allneighbors(dim):
    """src/brick-mpi.cpp:9-23: per axis "+", "none", "-"; axis 1 outermost."""
    out = []

    def rec(cur, idx):
        if idx > dim:
            out.append(cur)
            return
        rec(cur ^ _bit(idx), idx + 1)
        rec(cur, idx + 1)
        rec(cur ^ _bit(-idx), idx + 1)

    rec(0, 1)
    return out


def skin_all(dim):
    """weak/main-4d.cpp:103-105: every direction set except the centre, in allneighbors order"""
    s = allneighbors(dim)
    del s[len(s) // 2]
    return s


def sweep_skips():
    """brick_func of weak/main-4d.cpp:266-276: `skip` of the ST_ITER = 8 sweeps between two exchanges
    (0, then 1..6, then GZ); brick_stencil sweeps the grid minus skip // TILE bricks per side"""
    return [0] + list(range(1, ST_ITER - 1)) + [GZ]


class DecompND(object):
    def __init__(self, dims, depth=GZ, bdim=BDIM, skinlist=None, page_align=True, pagesize=4096, elem_bytes=8):
        dim = len(dims)
        self.dim = dim
        self.n = [int(x) // bdim for x in dims]          # interior extent in bricks
        self.g = [depth // bdim] * dim                    # ghost depth in bricks
        self.t = [n + 2 * g for n, g in zip(self.n, self.g)]
        self.skinlist = list(skin_all(dim) if skinlist is None else skinlist)
        # grid[.., x1, x0]: numpy axis order is outermost first
        self.grid = np.zeros(self.t[::-1], np.uint32)
        bsize = bdim ** dim * elem_bytes
        assert max(bsize, pagesize) % min(bsize, pagesize) == 0
        reserved = pagesize // bsize if bsize < pagesize else 1
        factor = reserved if page_align else 1
        self.pos = reserved

        def round_up():
            self.pos = (self.pos + factor - 1) // factor * factor

        def calc_pad(region):
            return factor - (self.region_size(region) + factor - 1) % factor - 1

        ns = len(self.skinlist)
        pad_first = []
        for i, s in enumerate(self.skinlist):
            prev = self.skinlist[i - 1] if i else 0
            nxt = self.skinlist[i + 1] if i + 1 < ns else 0
            pad_first.append(bin(prev & s).count("1") < bin(s & nxt).count("1"))

        self._number(0, 0)
        round_up()
        st_pos = [self.pos]
        self.sep_pos = [self.pos, 0, 0]
        self.skin_size = []
        for i, s in enumerate(self.skinlist):
            before = self.pos
            if pad_first[i]:
                self.pos += calc_pad(s)
            if s != 0:
                self._number(0, s)
                round_up()
            st_pos.append(self.pos)
            self.skin_size.append(self.pos - before)
        self.sep_pos[1] = self.pos

        self.ghost, self.skin = [], []
        for n in allneighbors(dim):
            if n == 0:
                continue
            mirrored = O.bitset_neg(n)
            last = -1
            g, sk = {}, {}
            for l in range(ns + 1):
                member = l < ns and (mirrored & self.skinlist[l]) == mirrored
                if member:
                    if last < 0:
                        last = l
                        fp = calc_pad(self.skinlist[l]) if pad_first[l] else 0
                        g = {"neighbor": n, "skin_st": l, "first_pad": fp, "pos": self.pos}
                        sk = {"neighbor": mirrored, "skin_st": l, "first_pad": fp, "pos": st_pos[l]}
                    if pad_first[l]:
                        self.pos += calc_pad(self.skinlist[l])
                    self._number(n, self.skinlist[l])
                    round_up()
                elif last >= 0:
                    lp = 0 if pad_first[l - 1] else calc_pad(self.skinlist[l - 1])
                    g.update(last_pad=lp, skin_ed=l, len=self.pos - g["pos"])
                    sk.update(last_pad=lp, skin_ed=l, len=st_pos[l] - sk["pos"])
                    self.ghost.append(g)
                    self.skin.append(sk)
                    last = -1
        self.sep_pos[2] = self.pos
        self.nbricks = self.pos
        self.adj = self._adjacency()

    def region_size(self, region):
        r = 1
        for d in range(1, self.dim + 1):
            r *= self.g[d - 1] if (has(region, d) or has(region, -d)) else self.n[d - 1] - 2 * self.g[d - 1]
        return r

    def _number(self, owner, region):
        """populate(owner, region, pos): the box of `region` inside the subdomain displaced by `owner`,
        numbered with axis 0 fastest"""
        sl = []
        for d in range(self.dim):
            n, g = self.n[d], self.g[d]
            if has(region, -(d + 1)):
                lo, cnt = g, g
            elif has(region, d + 1):
                lo, cnt = n, g
            else:
                lo, cnt = 2 * g, n - 2 * g
            if has(owner, d + 1):
                lo += n
            if has(owner, -(d + 1)):
                lo -= n
            if cnt <= 0:
                return
            sl.append(slice(lo, lo + cnt))
        shape = [s.stop - s.start for s in sl][::-1]
        count = int(np.prod(shape))
        self.grid[tuple(sl[::-1])] = (self.pos + np.arange(count, dtype=np.uint32)).reshape(shape)
        self.pos += count

    def _adjacency(self):
        """adj_populate: slot digits outermost axis first; outside the grid -> brick 0.  Rows of slots
        that no grid cell names (reserved / padding) stay zero."""
        dim, t = self.dim, self.t
        padded = np.zeros([x + 2 for x in t[::-1]], np.uint32)
        padded[tuple(slice(1, 1 + x) for x in t[::-1])] = self.grid
        adj = np.zeros((self.nbricks, 3 ** dim), np.uint32)
        flat = self.grid.reshape(-1)
        for slot in range(3 ** dim):
            rem, sl = slot, []
            for d in range(dim):  # least significant digit = axis 0
                off = rem % 3 - 1
                rem //= 3
                sl.append(slice(1 + off, 1 + off + t[d]))
            adj[flat, slot] = padded[tuple(sl[::-1])].reshape(-1)
        return adj

    def regions(self, name):
        r = self.ghost if name == "ghost" else self.skin
        keys = ("neighbor", "skin_st", "skin_ed", "pos", "len", "first_pad", "last_pad")
        return {k: [int(x[k]) for x in r] for k in keys}


def rank_map(procdims, coo):
    """populate(cart, decomp, 0, 1, coo) (include/brick-mpi.h:730-753): axis d of the set moves Cartesian
    coordinate coo[dim - d], "+d" to coordinate - 1 and "-d" to + 1; ranks row-major, periodic."""
    dim = len(procdims)
    out = {}
    for s in allneighbors(dim):
        c = list(coo)
        for d in range(1, dim + 1):
            if has(s, d):
                c[dim - d] -= 1
            if has(s, -d):
                c[dim - d] += 1
        r = 0
        for a in range(dim):
            r = r * procdims[a] + c[a] % procdims[a]
        out[s] = r
    return out


def exchange(storages, dec, procdims, step=ELEMS):
    """BrickDecomp::exchange (include/brick-mpi.h:466-495) on P virtual ranks, in place; all sends read
    the pre-exchange contents (skin ranges and ghost ranges are disjoint)."""
    P = len(storages)
    dim = dec.dim
    for r in range(P):
        coo, rr = [], r
        for n in reversed(procdims):
            coo.append(rr % n)
            rr //= n
        rm = rank_map(procdims, coo[::-1])
        for g, s in zip(dec.ghost, dec.skin):
            src = rm[g["neighbor"]]
            n = (g["len"] - g["first_pad"] - g["last_pad"]) * step
            ro = (g["pos"] + g["first_pad"]) * step
            so = (s["pos"] + s["first_pad"]) * step
            storages[r][ro:ro + n] = storages[src][so:so + n]
    assert dim == len(procdims) and P == int(np.prod(procdims))


# ---------------------------------------------------------------- C restatement of the stencil ----
def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _lib():
    L = O.lib()
    if not getattr(L, "_d4_ready", False):
        L.oracle_brick_sweep_4d.argtypes = [C.c_void_p] * 2 + [C.c_long] + [C.c_void_p] * 2 + [C.c_long] * 2
        L.oracle_sweep_list_4d.argtypes = [C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p]
        L.oracle_sweep_list_4d.restype = C.c_long
        L.oracle_array_sweep_4d.argtypes = [C.c_void_p] * 5
        L.oracle_copy_brick_4d.argtypes = [C.c_int] + [C.c_void_p] * 6 + [C.c_long]
        L._d4_ready = True
    return L


def brick_sweep(adj, blist, inp, out, in_step=ELEMS, out_step=ELEMS):
    adj = np.ascontiguousarray(adj, np.uint32)
    blist = np.ascontiguousarray(blist, np.uint32)
    assert adj.shape[1] == NADJ and inp.dtype == np.float64 and out.dtype == np.float64
    _lib().oracle_brick_sweep_4d(_p(adj), _p(blist), len(blist), _p(inp), _p(out), in_step, out_step)


def sweep_list(grid, s):
    """bricks of one sweep of weak/main-4d.cpp:62-75, `s` bricks inside the grid on each side"""
    g = np.ascontiguousarray(grid, np.uint32)
    gd = np.asarray(g.shape[::-1], np.uint32)
    L = _lib()
    n = L.oracle_sweep_list_4d(_p(g), _p(gd), s, None)
    out = np.zeros(max(n, 1), np.uint32)
    L.oracle_sweep_list_4d(_p(g), _p(gd), s, _p(out))
    return out[:n]


def array_sweep(inp, out, lo, hi):
    """inp/out: 4-D float64 arrays indexed [l][k][j][i]; box [lo, hi) per axis, contiguous axis first"""
    ext = np.asarray(inp.shape[::-1], np.int64)
    lo, hi = np.asarray(lo, np.int64), np.asarray(hi, np.int64)
    _lib().oracle_array_sweep_4d(_p(inp), _p(out), _p(ext), _p(lo), _p(hi))


def copy_brick(to_brick, dim, pad, ghost, arr, grid, bdat, step=ELEMS):
    d, p, g = (np.asarray(x, np.int64) for x in (dim, pad, ghost))
    grid = np.ascontiguousarray(grid, np.uint32)
    _lib().oracle_copy_brick_4d(1 if to_brick else 0, _p(d), _p(p), _p(g), _p(arr), _p(grid), _p(bdat), step)


# ---------------------------------------------------------------- the compiled reference ----
_ref = {}


def ref_available():
    return all(os.path.exists(os.path.join(HERE, "_ref", n)) for n in ("libref4d.so", "libref4d_unalign.so"))


def _reflib(unaligned=False):
    name = "libref4d_unalign.so" if unaligned else "libref4d.so"
    if name not in _ref:
        _ref[name] = C.CDLL(os.path.join(HERE, "_ref", name))
    return _ref[name]


class Ref4D(object):
    """BrickDecomp<4,4,4,4,4>(dims, 8).initialize(allneighbors minus centre) of the compiled reference"""

    def __init__(self, dims, gz=GZ, page_align=True):
        L = _reflib(not page_align)
        pre = "ref4a_" if page_align else "ref4u_"
        f = lambda n: getattr(L, pre + n)  # noqa: E731
        f("new").restype = C.c_void_p
        f("new").argtypes = [C.c_void_p, C.c_uint]
        for n in ("free", "nbricks", "sep_pos", "num_regions", "num_skin", "skin_size", "grid", "adj"):
            f(n).argtypes = [C.c_void_p] * (1 if n in ("free", "nbricks", "num_regions", "num_skin") else 2)
        f("regions").argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 7
        f("rank_map").argtypes = [C.c_void_p] * 5
        d = np.asarray(dims, np.uint32)
        self._f, self._h = f, f("new")(_p(d), gz)
        h = self._h
        self.nbricks = f("nbricks")(h)
        self.sep_pos = np.zeros(3, np.uint32)
        f("sep_pos")(h, _p(self.sep_pos))
        t = [int(x) // 4 + 2 * (gz // 4) for x in dims]
        self.grid = np.zeros(t[::-1], np.uint32)
        f("grid")(h, _p(self.grid))
        self.adj = np.zeros((self.nbricks, NADJ), np.uint32)
        f("adj")(h, _p(self.adj))
        nr, ns = f("num_regions")(h), f("num_skin")(h)
        self.skin_size = np.zeros(ns, np.int64)
        f("skin_size")(h, _p(self.skin_size))
        self.regions = {}
        for which, name in ((0, "ghost"), (1, "skin")):
            nb = np.zeros(nr, np.uint64)
            cols = [np.zeros(nr, np.uint32) for _ in range(6)]
            f("regions")(h, which, _p(nb), *[_p(c) for c in cols])
            self.regions[name] = dict(zip(("neighbor", "skin_st", "skin_ed", "pos", "len", "first_pad", "last_pad"),
                                          [nb] + cols))

    def rank_map(self, procdims, coo):
        keys, ranks = np.zeros(NADJ, np.uint64), np.zeros(NADJ, np.int32)
        pd, cc = np.asarray(procdims, np.int32), np.asarray(coo, np.int32)
        self._f("rank_map")(self._h, _p(pd), _p(cc), _p(keys), _p(ranks))
        return {int(k): int(r) for k, r in zip(keys, ranks)}

    def close(self):
        if self._h:
            self._f("free")(self._h)
            self._h = None


def ref_brick_sweep(adj, blist, inp, out, step=ELEMS):
    """one sweep of the reference's generated kernel over row-major bricks"""
    L = _reflib()
    L.ref4d_brick_sweep_rm.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_long]
    adj = np.ascontiguousarray(adj, np.uint32)
    blist = np.ascontiguousarray(blist, np.uint32)
    L.ref4d_brick_sweep_rm(_p(adj), adj.shape[0], _p(blist), len(blist), _p(inp), _p(out), step)


def ref_brick_sweep_timed(adj, blist, a, b, step=ELEMS, reps=4):
    L = _reflib()
    L.ref4d_brick_sweep_timed.restype = C.c_double
    L.ref4d_brick_sweep_timed.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p,
                                          C.c_long, C.c_int]
    adj = np.ascontiguousarray(adj, np.uint32)
    blist = np.ascontiguousarray(blist, np.uint32)
    return L.ref4d_brick_sweep_timed(_p(adj), adj.shape[0], _p(blist), len(blist), _p(a), _p(b), step, reps)
