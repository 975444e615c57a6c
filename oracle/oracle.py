"""oracle/oracle.py -- TEST INFRASTRUCTURE, not product code.

ctypes front end of oracle/_build/liboracle.so (stencil_oracle.c + decomp_oracle.c) plus the
numpy restatement of the halo exchange and of the process-grid helpers.  It is the checker
for the CUDA path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import it; the product (bricklib_b200/) never does.

PINNED against the compiled reference: see tests/test_oracle_golden.py + oracle/gen_golden.py.

Exchange restated from include/brick-mpi.h:466-495 (BrickDecomp::exchange): for every region
i, rank r receives (ghost[i].len - pads)*step elements at dat[(ghost[i].pos+first_pad)*step]
from rank_map_r[ghost[i].neighbor]; that peer sends the same count from
dat[(skin[i].pos+first_pad)*step] (tag i pairs them).  Process grid from weak/args.cpp:105-115
(MPI_Dims_create + periodic MPI_Cart_create; ranks row-major per the MPI standard).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIBPATH = os.path.join(HERE, "_build", "liboracle.so")

SKIN3D_GOOD = [  # src/brick-mpi.cpp:25-52
    (1,), (1, -3), (1, 2, -3), (1, 2), (1, 2, 3), (2, 3), (2,), (2, -3), (-1, 2, -3), (-1, 2),
    (-1, 2, 3), (-1, 3), (-1,), (-3,), (-1, -3), (-1, -2, -3), (-1, -2), (-1, -2, 3), (-2, 3),
    (-2,), (-2, -3), (1, -2, -3), (1, -2), (1, -2, 3), (1, 3), (3,)]


def bitset(elems):
    s = 0
    for d in elems:
        s ^= 1 << (31 - d if d < 0 else d)
    return s


def bitset_neg(s):
    return ((s & ((1 << 32) - 1)) << 31) | (s >> 31)


class _Region(C.Structure):
    _fields_ = [("neighbor", C.c_uint64), ("skin_st", C.c_uint), ("skin_ed", C.c_uint),
                ("pos", C.c_uint), ("len", C.c_uint), ("first_pad", C.c_uint), ("last_pad", C.c_uint)]


_REGION_DT = np.dtype([("neighbor", np.uint64), ("skin_st", np.uint32), ("skin_ed", np.uint32),
                       ("pos", np.uint32), ("len", np.uint32), ("first_pad", np.uint32),
                       ("last_pad", np.uint32)], align=True)

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIBPATH):
            build()
        L = C.CDLL(LIBPATH)
        vp, u32p, f64p = C.c_void_p, C.POINTER(C.c_uint), C.POINTER(C.c_double)
        L.oracle_points.restype = C.c_int
        L.oracle_points.argtypes = [C.c_int, vp, vp, vp]
        L.oracle_brick_sweep.restype = C.c_int
        L.oracle_brick_sweep.argtypes = [C.c_int, vp, vp, C.c_long, vp, C.c_long, vp, C.c_long, vp]
        L.oracle_brick_sweep_f32.restype = C.c_int
        L.oracle_brick_sweep_f32.argtypes = [C.c_int, vp, vp, C.c_long, vp, C.c_long, vp, C.c_long, vp]
        L.oracle_array_sweep.restype = C.c_int
        L.oracle_array_sweep.argtypes = [C.c_int, vp, vp] + [C.c_long] * 9 + [vp]
        L.oracle_copy_brick.argtypes = [C.c_int, vp, vp, vp, vp, vp, vp, C.c_long]
        L.oracle_decomp_build.restype = vp
        L.oracle_decomp_build.argtypes = [vp, C.c_uint, vp, C.c_int, C.c_int, C.c_long, C.c_long]
        L.oracle_decomp_free.argtypes = [vp]
        L.oracle_decomp_nbricks.restype = C.c_uint
        L.oracle_decomp_nbricks.argtypes = [vp]
        L.oracle_decomp_nregions.restype = C.c_int
        L.oracle_decomp_nregions.argtypes = [vp]
        L.oracle_decomp_get.argtypes = [vp] * 7
        L.oracle_rank_map.argtypes = [vp] * 4
        L.oracle_init_grid.argtypes = [vp] * 3
        L.oracle_allneighbors.restype = C.c_int
        L.oracle_allneighbors.argtypes = [vp]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def points(which, coeff=None):
    off = np.zeros((125, 3), np.int32)
    c = np.zeros(125, np.float64)
    cin = None if coeff is None else np.ascontiguousarray(coeff, np.float64)
    n = lib().oracle_points(which, _p(cin), _p(off), _p(c))
    if n < 0:
        raise ValueError("unknown stencil %r" % (which,))
    return off[:n].copy(), c[:n].copy()


def brick_sweep(which, adj, blist, inp, out, in_step=512, out_step=512, coeff=None):
    """out[b] = stencil(inp) for b in blist; inp/out are 1-D float64 views starting at the
    field offset (row-major 8x8x8 bricks)."""
    adj = np.ascontiguousarray(adj, np.uint32)
    blist = np.ascontiguousarray(blist, np.uint32)
    cin = None if coeff is None else np.ascontiguousarray(coeff, np.float64)
    assert inp.dtype == np.float64 and out.dtype == np.float64
    r = lib().oracle_brick_sweep(which, _p(adj), _p(blist), blist.size, _p(inp), in_step, _p(out),
                                 out_step, _p(cin))
    if r:
        raise ValueError("unknown stencil %r" % (which,))
    return out


def brick_sweep_f32(which, adj, blist, inp, out, in_step=512, out_step=512, coeff=None):
    """fp32 bricks: out[b] = stencil(inp) computed in float (1-D float32 views, row-major bricks)."""
    adj = np.ascontiguousarray(adj, np.uint32)
    blist = np.ascontiguousarray(blist, np.uint32)
    cin = None if coeff is None else np.ascontiguousarray(coeff, np.float64)
    assert inp.dtype == np.float32 and out.dtype == np.float32
    r = lib().oracle_brick_sweep_f32(which, _p(adj), _p(blist), blist.size, _p(inp), in_step, _p(out),
                                     out_step, _p(cin))
    if r:
        raise ValueError("unknown stencil %r for fp32" % (which,))
    return out


def array_sweep(which, inp, out, box, coeff=None):
    s2, s1, s0 = inp.shape
    cin = None if coeff is None else np.ascontiguousarray(coeff, np.float64)
    r = lib().oracle_array_sweep(which, _p(inp), _p(out), s0, s1, s2, *[int(x) for x in box], _p(cin))
    if r:
        raise ValueError("unknown stencil %r" % (which,))
    return out


def copy_brick(to_brick, dim, pad, ghost, arr, grid, bdat, step=512):
    d, p, g = (np.asarray(x, np.int64) for x in (dim, pad, ghost))
    grid = np.ascontiguousarray(grid, np.uint32)
    lib().oracle_copy_brick(1 if to_brick else 0, _p(d), _p(p), _p(g), _p(arr), _p(grid), _p(bdat), step)


def init_grid(dimlist):
    d = np.asarray(dimlist, np.int64)
    n = int(d[0] * d[1] * d[2])
    grid = np.zeros(n, np.uint32)
    adj = np.zeros((n, 27), np.uint32)
    lib().oracle_init_grid(_p(d), _p(grid), _p(adj))
    return grid.reshape(int(d[2]), int(d[1]), int(d[0])), adj


def allneighbors():
    out = np.zeros(27, np.uint64)
    lib().oracle_allneighbors(_p(out))
    return [int(x) for x in out]


class Decomp(object):
    """Restated BrickDecomp<3,8,8,8>(dims, depth).initialize(skinlist)."""

    def __init__(self, dims, depth=8, skinlist=None, page_align=False, pagesize=4096, brick_bytes=4096):
        L = lib()
        skinlist = SKIN3D_GOOD if skinlist is None else skinlist
        sk = np.array([bitset(s) if not isinstance(s, int) else s for s in skinlist], np.uint64)
        d = np.asarray(dims, np.uint32)
        h = L.oracle_decomp_build(_p(d), depth, _p(sk), len(sk), int(page_align), pagesize, brick_bytes)
        self.dims = tuple(int(x) for x in dims)
        self.depth = depth
        self.skinlist = [int(x) for x in sk]
        self.nbricks = L.oracle_decomp_nbricks(h)
        nr = L.oracle_decomp_nregions(h)
        t = [x // 8 + 2 * (depth // 8) for x in self.dims]
        self.tdims = t
        self.sep_pos = np.zeros(3, np.uint32)
        grid = np.zeros(t[0] * t[1] * t[2], np.uint32)
        self.adj = np.zeros((self.nbricks, 27), np.uint32)
        ghost = np.zeros(nr, _REGION_DT)
        skin = np.zeros(nr, _REGION_DT)
        assert _REGION_DT.itemsize == C.sizeof(_Region)
        self.skin_size = np.zeros(len(sk), np.int64)
        L.oracle_decomp_get(h, _p(self.sep_pos), _p(grid), _p(self.adj), _p(ghost), _p(skin),
                            _p(self.skin_size))
        L.oracle_decomp_free(h)
        self.grid = grid.reshape(t[2], t[1], t[0])
        self.ghost = ghost
        self.skin = skin

    def regions(self, name):
        r = self.ghost if name == "ghost" else self.skin
        return {k: r[k].copy() for k in r.dtype.names}


def rank_map(procdims, coo):
    keys = np.zeros(27, np.uint64)
    ranks = np.zeros(27, np.int32)
    pd, cc = np.asarray(procdims, np.int32), np.asarray(coo, np.int32)
    lib().oracle_rank_map(_p(pd), _p(cc), _p(keys), _p(ranks))
    return {int(k): int(r) for k, r in zip(keys, ranks)}


def dims_create(nnodes, ndims=3):
    """MPI_Dims_create(nnodes, ndims, {0..}): balanced factors, non-increasing
    (8 -> 2x2x2, 4 -> 2x2x1, 2 -> 2x1x1, 6 -> 3x2x1, 1 -> 1x1x1)."""
    primes, n, p = [], nnodes, 2
    while n > 1:
        while n % p == 0:
            primes.append(p)
            n //= p
        p += 1
    dims = [1] * ndims
    for f in sorted(primes, reverse=True):
        dims[dims.index(min(dims))] *= f
    return sorted(dims, reverse=True)


def cart_coords(rank, procdims):
    coo = []
    for n in reversed(procdims):
        coo.append(rank % n)
        rank //= n
    return list(reversed(coo))


def exchange(storages, dec, procdims, step=512):
    """In-place emulation of BrickDecomp::exchange on P virtual ranks (row-major ranks on a
    periodic procdims grid).  storages[r] is rank r's 1-D float64 brick storage."""
    P = len(storages)
    snap = [s.copy() for s in storages]  # MPI semantics: every send buffer is the pre-exchange skin
    for r in range(P):
        rm = rank_map(procdims, cart_coords(r, procdims))
        for i in range(len(dec.ghost)):
            g, s = dec.ghost[i], dec.skin[i]
            src = rm[int(g["neighbor"])]
            n = (int(g["len"]) - int(g["first_pad"]) - int(g["last_pad"])) * step
            d0 = (int(g["pos"]) + int(g["first_pad"])) * step
            s0 = (int(s["pos"]) + int(s["first_pad"])) * step
            storages[r][d0:d0 + n] = snap[src][s0:s0 + n]
    return storages


def multistage_packages(dec):
    """Restated BrickDecomp::multiStageExchangeView (include/brick-mpi.h:587-681): the SHIFT exchange.
    Stage d = 1..3 exchanges along axis d only; what a corner or edge neighbour owns travels in
    several hops through the ghost zones of the face neighbours.  Returns, per stage, a list of
    (n, recv_pieces, send_pieces) in allneighbors order, n = the direction received from (the matching
    send goes to the mirrored direction); pieces are (first brick, brick count), in packing order.
    Receive side: the ghost regions owned by n, then every not yet received ghost region whose owner
    is n plus already exchanged axes.  Send side: the skin regions facing the mirrored direction, then,
    for each of those further ghost regions and each skin section it spans, the same section inside
    this rank's own ghost copy of the owner-minus-n neighbour (received in an earlier stage)."""
    ghost, skin, ssz = dec.ghost, dec.skin, [int(x) for x in dec.skin_size]
    stages, done = [], 0
    for d in (1, 2, 3):
        axis = bitset((d,)) | bitset((-d,))
        pkgs = []
        for n in allneighbors():
            if n == 0 or (n & ~axis):
                continue
            recv = [(int(g["pos"]), int(g["len"])) for g in ghost if int(g["neighbor"]) == n]
            mirrored = bitset_neg(n)
            send = [(int(k["pos"]), int(k["len"])) for k in skin if int(k["neighbor"]) == mirrored]
            reach = n | done
            for g in ghost:
                owner = int(g["neighbor"])
                if owner == n or (owner & ~reach) or not (owner & ~done):
                    continue
                recv.append((int(g["pos"]), int(g["len"])))
                via = owner ^ n
                for sec in range(int(g["skin_st"]), int(g["skin_ed"])):
                    for g2 in ghost:
                        if int(g2["neighbor"]) == via and int(g2["skin_st"]) <= sec < int(g2["skin_ed"]):
                            first = int(g2["pos"]) + sum(ssz[int(g2["skin_st"]):sec])
                            send.append((first, ssz[sec]))
            pkgs.append((n, recv, send))
        stages.append(pkgs)
        done |= axis
    return stages


def multistage_exchange(storages, dec, procdims, step=512):
    """In-place emulation of MultiStageExchangeView::exchange (brick-mpi.h:126-176) on P virtual
    ranks: per stage every rank packs its send pieces, the package for direction n goes to the rank
    in the mirrored direction, and is unpacked over that rank's receive pieces for n."""
    P = len(storages)
    maps = [rank_map(procdims, cart_coords(r, procdims)) for r in range(P)]
    for pkgs in multistage_packages(dec):
        snap = [s.copy() for s in storages]
        for r in range(P):
            for n, recv, _ in pkgs:
                src = maps[r][n]
                sent = [sp for (m, _, sp) in pkgs if m == n][0]
                packed = np.concatenate([snap[src][a * step:(a + l) * step] for a, l in sent])
                assert maps[src][bitset_neg(n)] == r
                at = 0
                for a, l in recv:
                    storages[r][a * step:(a + l) * step] = packed[at:at + l * step]
                    at += l * step
                assert at == packed.size
    return storages


# ---------------------------------------------------------------------------
# strong scaling: restatement of strong/args.cpp:36-56,104-113 (getrank, ownership split) and of the
# link construction of strong/main.cu:188-247 -- ghost[i] of subdomain T is skin[i] of the periodic
# neighbour of T in direction ghost[i].neighbor.  Morton coordinate d <-> axis d+1 (zmort.h:18-105).
def morton3(ci, cj, ck):
    """ZMORT()[ck][cj][ci]: bit b of ci lands on bit 3b, cj on 3b+1, ck on 3b+2."""
    z = 0
    for b in range(21):
        z |= ((ci >> b) & 1) << (3 * b) | ((cj >> b) & 1) << (3 * b + 1) | ((ck >> b) & 1) << (3 * b + 2)
    return z


def morton3_coords(z):
    c = [0, 0, 0]
    for b in range(21):
        for d in range(3):
            c[d] |= ((z >> (3 * b + d)) & 1) << b
    return tuple(c)


def strong_owner(sid, total, nranks):
    base, extra = divmod(total, nranks)
    split = extra * (base + 1)
    return sid // (base + 1) if sid < split else extra + (sid - split) // base


def strong_range(rank, total, nranks):
    base, extra = divmod(total, nranks)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def strong_neighbor(sid, neighbor_set, subdim):
    c = list(morton3_coords(sid))
    for d in range(3):
        if neighbor_set & (1 << (d + 1)):
            c[d] = (c[d] + 1) % subdim
        elif neighbor_set & (1 << (31 + d + 1)):
            c[d] = (c[d] - 1) % subdim
    return morton3(*c)


def strong_exchange(storages, dec, subdim, step=512):
    """storages: {morton id: 1-D float64 storage}.  In-place ghost fill of every subdomain."""
    snap = {k: v.copy() for k, v in storages.items()}
    for sid, st in storages.items():
        for i in range(len(dec.ghost)):
            g, s = dec.ghost[i], dec.skin[i]
            src = strong_neighbor(sid, int(g["neighbor"]), subdim)
            n = (int(g["len"]) - int(g["first_pad"]) - int(g["last_pad"])) * step
            d0 = (int(g["pos"]) + int(g["first_pad"])) * step
            s0 = (int(s["pos"]) + int(s["first_pad"])) * step
            st[d0:d0 + n] = snap[src][s0:s0 + n]
    return storages
