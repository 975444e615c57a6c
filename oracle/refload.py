"""oracle/refload.py -- TEST INFRASTRUCTURE, not product code.

ctypes loader for oracle/_ref/libref_<isa>.so: the UNMODIFIED reference CPU path
(codegen brick kernels, ST_CPU array stencils, BrickDecomp::initialize/populate)
compiled by oracle/Makefile from /root/reference.  Importable only from tests/,
oracle/gen_golden.py, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Never touches /root/reference at run time.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFDIR = os.path.join(HERE, "_ref")

_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")

STENCILS = (0, 1, 7, 13, 25, 125)
STENCILS_F32 = (7, 13, 25, 125)  # built with -DbElem=float (AVX-512 only): symbols reff<N>_*  # 0 = stencils/7pt.py with coeff[], 1 = cond.py (same coeff[]); others = mpi<N>pt.py


def host_isa():
    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = line
                    break
    except OSError:
        pass
    fl = set(flags.split())
    need512 = {"avx512f", "avx512cd", "avx512bw", "avx512dq", "avx512vl"}
    if need512 <= fl:
        return "avx512"
    if "avx2" in fl and "fma" in fl:
        return "avx2"
    return None


_lib = None


def available():
    isa = host_isa()
    return isa is not None and os.path.exists(os.path.join(REFDIR, "libref_%s.so" % isa))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    isa = host_isa()
    if isa is None:
        raise RuntimeError("oracle/_ref needs an AVX2 or AVX-512 host")
    path = os.path.join(REFDIR, "libref_%s.so" % isa)
    L = C.CDLL(path)
    for w in STENCILS:
        p = "ref%d_" % w
        getattr(L, p + "st_iter").restype = C.c_int
        getattr(L, p + "vsvec").restype = C.c_char_p
        getattr(L, p + "max_threads").restype = C.c_int
        getattr(L, p + "fold_table").argtypes = [_u32p]
        sw = [_u32p, C.c_uint, _u32p, C.c_long, _f64p, _f64p, C.c_long, C.c_void_p]
        getattr(L, p + "brick_sweep").argtypes = sw
        getattr(L, p + "brick_sweep_rm").argtypes = sw
        getattr(L, p + "brick_sweep_timed").argtypes = sw + [C.c_int]
        getattr(L, p + "brick_sweep_timed").restype = C.c_double
        getattr(L, p + "array_sweep").argtypes = [_f64p, _f64p] + [C.c_long] * 9 + [C.c_void_p]
    L.ref_decomp_new.restype = C.c_void_p
    L.ref_decomp_new.argtypes = [_u32p, C.c_uint, C.c_int]
    L.ref_decomp_free.argtypes = [C.c_void_p]
    L.ref_decomp_nbricks.restype = C.c_uint
    L.ref_decomp_nbricks.argtypes = [C.c_void_p]
    L.ref_decomp_sep_pos.argtypes = [C.c_void_p, _u32p]
    L.ref_decomp_num_regions.restype = C.c_uint
    L.ref_decomp_num_regions.argtypes = [C.c_void_p]
    L.ref_decomp_regions.argtypes = [C.c_void_p, C.c_int, _u64p] + [_u32p] * 6
    L.ref_decomp_num_skin.restype = C.c_uint
    L.ref_decomp_num_skin.argtypes = [C.c_void_p]
    L.ref_decomp_skin_size.argtypes = [C.c_void_p, _i64p]
    L.ref_decomp_grid.argtypes = [C.c_void_p, _u32p]
    L.ref_decomp_adj.argtypes = [C.c_void_p, _u32p]
    L.ref_decomp_rank_map.argtypes = [C.c_void_p, _i32p, _i32p, _u64p, _i32p]
    L._path = path
    _lib = L
    return L


def aligned_empty(n, dtype=np.float64, align=2048):
    """The reference allocates bricks with ALIGN = 2048 (include/brick.h:15) and its AVX
    kernels use aligned vector loads; numpy only guarantees 16/64 bytes."""
    item = np.dtype(dtype).itemsize
    raw = np.empty(n * item + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * item].view(dtype)


def _aligned(a, align=64):
    a = np.ascontiguousarray(a)
    if a.ctypes.data % align == 0:
        return a
    b = aligned_empty(a.size, a.dtype).reshape(a.shape)
    b[...] = a
    return b


def _coeff_ptr(coeff):
    if coeff is None:
        return None, None
    c = np.ascontiguousarray(coeff, dtype=np.float64)
    return c, c.ctypes.data_as(C.c_void_p)


def brick_sweep_rm(which, adj, blist, inp, out, step, coeff=None):
    """One reference codegen sweep over row-major brick storages (in place on `out`)."""
    L = lib()
    keep, cp = _coeff_ptr(coeff)
    adj = np.ascontiguousarray(adj, dtype=np.uint32)
    blist = np.ascontiguousarray(blist, dtype=np.uint32)
    ai, ao = _aligned(inp.reshape(-1)), _aligned(out.reshape(-1))
    getattr(L, "ref%d_brick_sweep_rm" % which)(adj.reshape(-1), adj.shape[0], blist, blist.size,
                                               ai, ao, step, cp)
    if ao is not out:
        out.reshape(-1)[...] = ao
    return out


def brick_sweep_timed(which, adj, blist, a, b, step, coeff=None, reps=4):
    L = lib()
    keep, cp = _coeff_ptr(coeff)
    adj = np.ascontiguousarray(adj, dtype=np.uint32)
    blist = np.ascontiguousarray(blist, dtype=np.uint32)
    assert a.ctypes.data % 64 == 0 and b.ctypes.data % 64 == 0, "use refload.aligned_empty"
    return getattr(L, "ref%d_brick_sweep_timed" % which)(adj.reshape(-1), adj.shape[0], blist,
                                                         blist.size, a.reshape(-1), b.reshape(-1),
                                                         step, cp, reps)


def array_sweep(which, inp, out, box, coeff=None):
    """Reference array stencil on arrays shaped [s2][s1][s0]; box = (k0,k1,j0,j1,i0,i1)."""
    L = lib()
    keep, cp = _coeff_ptr(coeff)
    s2, s1, s0 = inp.shape
    getattr(L, "ref%d_array_sweep" % which)(inp.reshape(-1), out.reshape(-1), s0, s1, s2,
                                            *[int(x) for x in box], cp)
    return out


def available_f32():
    """the fp32 build of the reference (bElem = float) exists for AVX-512 hosts only"""
    if not available() or host_isa() != "avx512":
        return False
    try:
        return hasattr(lib(), "reff125_brick_sweep_rm")
    except OSError:
        return False


def brick_sweep_rm_f32(which, adj, blist, inp, out, step):
    """One sweep of the reference's generated brick kernel built with bElem = float over row-major
    float32 brick storages (in place on `out`)."""
    L = lib()
    fn = getattr(L, "reff%d_brick_sweep_rm" % which)
    fn.argtypes = [_u32p, C.c_uint, _u32p, C.c_long, C.c_void_p, C.c_void_p, C.c_long, C.c_void_p]
    fn.restype = None
    adj = np.ascontiguousarray(adj, dtype=np.uint32)
    blist = np.ascontiguousarray(blist, dtype=np.uint32)
    assert inp.dtype == np.float32 and out.dtype == np.float32
    ai = aligned_empty(inp.size, np.float32)
    ai[:] = inp.reshape(-1)
    ao = aligned_empty(out.size, np.float32)
    ao[:] = out.reshape(-1)
    fn(adj.reshape(-1), adj.shape[0], blist, blist.size, ai.ctypes.data, ao.ctypes.data, step, None)
    out.reshape(-1)[...] = ao
    return out


def max_threads():
    return lib().ref7_max_threads()


def vsvec():
    return lib().ref7_vsvec().decode()


class RefDecomp(object):
    """Tables of the reference BrickDecomp<3,8,8,8>(dims, gz).initialize(skin)."""

    def __init__(self, dims, gz=8, skin=0):
        L = lib()
        d = np.asarray(dims, dtype=np.uint32)
        h = L.ref_decomp_new(d, gz, skin)
        try:
            self.dims = tuple(int(x) for x in dims)
            self.gz = gz
            self.nbricks = L.ref_decomp_nbricks(h)
            sp = np.zeros(3, np.uint32)
            L.ref_decomp_sep_pos(h, sp)
            self.sep_pos = sp
            n = L.ref_decomp_num_regions(h)
            self.regions = {}
            for which, name in ((0, "ghost"), (1, "skin")):
                nb = np.zeros(n, np.uint64)
                cols = [np.zeros(n, np.uint32) for _ in range(6)]
                L.ref_decomp_regions(h, which, nb, *cols)
                self.regions[name] = dict(neighbor=nb, skin_st=cols[0], skin_ed=cols[1], pos=cols[2],
                                          len=cols[3], first_pad=cols[4], last_pad=cols[5])
            ns = L.ref_decomp_num_skin(h)
            ss = np.zeros(ns, np.int64)
            L.ref_decomp_skin_size(h, ss)
            self.skin_size = ss
            t = [x // 8 + 2 * (gz // 8) for x in self.dims]
            g = np.zeros(t[0] * t[1] * t[2], np.uint32)
            L.ref_decomp_grid(h, g)
            self.grid = g.reshape(t[2], t[1], t[0])
            a = np.zeros(self.nbricks * 27, np.uint32)
            L.ref_decomp_adj(h, a)
            self.adj = a.reshape(self.nbricks, 27)
            self.adj[0, :] = 0  # row 0 (the sentinel brick) is never written by the reference
            self._h = h
            self._L = L
        except Exception:
            L.ref_decomp_free(h)
            raise

    def rank_map(self, procdims, coo):
        keys = np.zeros(27, np.uint64)
        ranks = np.zeros(27, np.int32)
        self._L.ref_decomp_rank_map(self._h, np.asarray(procdims, np.int32), np.asarray(coo, np.int32),
                                    keys, ranks)
        return {int(k): int(r) for k, r in zip(keys, ranks)}

    def close(self):
        if self._h:
            self._L.ref_decomp_free(self._h)
            self._h = None
