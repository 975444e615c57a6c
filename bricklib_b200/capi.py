"""ctypes binding of the C ABI in include/bricks_b200.h (libbricks_b200.so).

This is the only way Python reaches the CUDA path: there is no CPU fallback.  Loading fails
loudly when the library has not been built (__graft_entry__.build()).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# BRICKS_LIB: load an experiment build (bricklib_b200/build.py BRICKS_EXP_NAME) instead
LIBPATH = os.environ.get("BRICKS_LIB") or os.path.join(HERE, "lib", "libbricks_b200.so")

ST_7PT_COEFF, ST_7PT, ST_13PT, ST_25PT, ST_125PT = 0, 7, 13, 25, 125
ST_COND = 1  # stencils/cond.py (d3cond): 7pt.py's coefficients on max(x, 0), |sum|
ST_27PT_COEFF = 27  # d3pt27 (stencils/3axis.cpp:165-247): 27 runtime coefficients
STREAM_COMPUTE, STREAM_COMM, STREAM_H2D, STREAM_D2H = 0, 1, 2, 3
#: ghost-zone sweeps per exchange, stencils/fake.h:39-344 (ST_ITER) -- GZ / radius
ST_ITER = {ST_7PT: 8, ST_13PT: 4, ST_25PT: 2, ST_125PT: 4, ST_7PT_COEFF: 8}
RADIUS = {ST_7PT: 1, ST_13PT: 2, ST_25PT: 4, ST_125PT: 2, ST_7PT_COEFF: 1, ST_27PT_COEFF: 1, ST_COND: 1}


class BricksError(RuntimeError):
    pass


class RegionLink(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("len", C.c_size_t)]


_lib = None

_vp, _vpp = C.c_void_p, C.POINTER(C.c_void_p)
_SIGS = {
    "bricks_last_error": (C.c_char_p, []),
    "bricks_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "bricks_ctx_create": (C.c_int, [C.c_int, _vpp]),
    "bricks_ctx_destroy": (C.c_int, [_vp]),
    "bricks_ctx_device": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "bricks_malloc": (C.c_int, [_vp, _vpp, C.c_size_t]),
    "bricks_free": (C.c_int, [_vp, _vp]),
    "bricks_memset": (C.c_int, [_vp, _vp, C.c_int, C.c_size_t]),
    "bricks_host_alloc": (C.c_int, [_vpp, C.c_size_t]),
    "bricks_host_alloc_local": (C.c_int, [_vp, _vpp, C.c_size_t, C.POINTER(C.c_int)]),
    "bricks_host_free": (C.c_int, [_vp]),
    "bricks_h2d": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int]),
    "bricks_d2h": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int]),
    "bricks_d2d": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_int]),
    "bricks_sync": (C.c_int, [_vp, C.c_int]),
    "bricks_stream_wait": (C.c_int, [_vp, C.c_int, C.c_int]),
    "bricks_stream_handle": (C.c_int, [_vp, C.c_int, _vpp]),
    "bricks_timer_start": (C.c_int, [_vp, C.c_int, C.c_int]),
    "bricks_timer_stop": (C.c_int, [_vp, C.c_int, C.c_int]),
    "bricks_timer_elapsed_ms": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float)]),
    "bricks_plan_create": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint, C.c_int, _vp, _vpp]),
    "bricks_plan_derive": (C.c_int, [_vp, C.c_int, _vp, _vpp]),
    "bricks_plan_destroy": (C.c_int, [_vp]),
    "bricks_plan_device_tables": (C.c_int, [_vp, _vpp, _vpp]),
    "bricks_stencil": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, C.c_size_t, C.c_size_t, _vp, C.c_int]),
    "bricks_stencil_naive": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, C.c_size_t, C.c_size_t, _vp, C.c_int]),
    "bricks_stencil_f32": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, C.c_size_t, C.c_size_t, _vp, C.c_int]),
    "bricks_stencil_naive_f32": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, C.c_size_t, C.c_size_t, _vp, C.c_int]),
    "bricks_plan4d_create": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint, _vpp]),
    "bricks_plan4d_destroy": (C.c_int, [_vp]),
    "bricks_stencil4d": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_size_t, C.c_uint, C.c_int]),
    "bricks_stencil4d_naive": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_size_t, C.c_uint, C.c_int]),
    "bricks_launch_count": (C.c_int, [C.POINTER(C.c_ulonglong)]),
    "bricks_links_upload": (C.c_int, [_vp, _vp, C.c_int, _vpp]),
    "bricks_copy_regions": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "bricks_links_free": (C.c_int, [_vp, _vp]),
    "bricks_copy_regions_static": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "bricks_copy_array": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, C.c_int]),
    "bricks_array_stencil7": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int]),
    "bricks_comm_unique_id": (C.c_int, [_vp]),
    "bricks_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "bricks_comm_init_config": (C.c_int, [_vp, C.c_int, C.c_int, _vp, C.c_int]),
    "bricks_comm_destroy": (C.c_int, [_vp]),
    "bricks_exchange_create": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vpp]),
    "bricks_exchange_destroy": (C.c_int, [_vp]),
    "bricks_exchange_run": (C.c_int, [_vp, _vp, C.c_int]),
    "bricks_decomp_step_create": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_size_t, _vpp, C.c_char_p, C.c_size_t]),
    "bricks_strong_step_create": (C.c_int, [_vp, _vp, C.c_int, _vp, _vp, C.c_size_t, _vpp, C.POINTER(C.c_size_t)]),
    "bricks_step_boxes": (C.c_int, [_vp, C.c_uint, C.c_int, _vp, C.POINTER(C.c_int)]),
    "bricks_step_create": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint, C.c_uint, _vpp]),
    "bricks_step_destroy": (C.c_int, [_vp]),
    "bricks_step_plan": (C.c_int, [_vp, C.c_int, _vpp]),
    "bricks_step_arena": (C.c_int, [_vp, C.c_size_t, _vpp]),
    "bricks_step_add_copy": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int]),
    "bricks_step_add_exchange": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "bricks_step_run": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, C.c_int]),
    "bricks_step_exchange": (C.c_int, [_vp, _vp]),
    "bricks_step_sweep": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, C.c_int]),
    "bricks_step_reset": (C.c_int, [_vp]),
    "bricks_step_p2p_export": (C.c_int, [_vp, _vp, C.c_int, _vp]),
    "bricks_step_p2p_connect": (C.c_int, [_vp, C.c_int, _vp]),
    "bricks_step_p2p_disable": (C.c_int, [_vp]),
    "bricks_step_trace": (C.c_int, [_vp, C.c_int]),
    "bricks_step_trace_read": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_int), _vp, _vp, _vp]),
    # include/bricks_b200_probe.h
    "bricks_probe_dfma": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_float)]),
    "bricks_probe_copy": (C.c_int, [_vp, C.c_size_t, C.c_int, C.POINTER(C.c_double)]),
    "bricks_probe_read": (C.c_int, [_vp, C.c_size_t, C.c_int, C.POINTER(C.c_double)]),
    "bricks_probe_mio": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "bricks_probe_tma_image": (C.c_int, [_vp, _vp, C.c_uint, C.c_size_t, C.c_uint, _vp]),
    "bricks_probe_tma_issue": (C.c_int, [_vp, _vp, C.c_uint, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_double),
                                          C.POINTER(C.c_double)]),
}


def _preload_nccl():
    """libbricks_b200.so needs libnccl.so.2.  When PyTorch's bundled NCCL wheel is installed, load
    THAT copy first so that this library and a later `import torch` share one NCCL (the system
    copy is older and lacks symbols libtorch_cuda wants)."""
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for base in (spec.submodule_search_locations if spec else []):
            path = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return
    except Exception:  # noqa: BLE001 -- fall back to the dynamic linker's search path
        pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIBPATH):
            raise BricksError("libbricks_b200.so is not built (run __graft_entry__.build()); "
                              "there is no CPU fallback")
        _preload_nccl()
        L = C.CDLL(LIBPATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise BricksError("bricks_b200 error %d: %s" % (rc, lib().bricks_last_error().decode()))


def _np_ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Context(object):
    """One device.  Mirrors the role of the gpu* macros of include/brick-cuda.h."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        check(lib().bricks_ctx_create(device, C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().bricks_ctx_destroy(self._h)
            self._h = C.c_void_p()

    # memory -----------------------------------------------------------------
    def malloc(self, nbytes):
        p = C.c_void_p()
        check(lib().bricks_malloc(self._h, C.byref(p), nbytes))
        return p.value

    def free(self, dptr):
        check(lib().bricks_free(self._h, C.c_void_p(dptr)))

    def memset(self, dptr, value, nbytes):
        check(lib().bricks_memset(self._h, C.c_void_p(dptr), value, nbytes))

    def h2d(self, dptr, host, stream=STREAM_COMPUTE, nbytes=None):
        host = np.ascontiguousarray(host)
        check(lib().bricks_h2d(self._h, C.c_void_p(dptr), _np_ptr(host), host.nbytes if nbytes is None else nbytes, stream))
        return host

    def d2h(self, host, dptr, stream=STREAM_COMPUTE, nbytes=None):
        assert host.flags["C_CONTIGUOUS"]
        check(lib().bricks_d2h(self._h, _np_ptr(host), C.c_void_p(dptr), host.nbytes if nbytes is None else nbytes, stream))
        return host

    def d2d(self, dst, src, nbytes, stream=STREAM_COMPUTE):
        check(lib().bricks_d2d(self._h, C.c_void_p(dst), C.c_void_p(src), nbytes, stream))

    def upload(self, host):
        """movBrickStorage / copyToDevice: allocate + copy, synchronous."""
        host = np.ascontiguousarray(host)
        d = self.malloc(host.nbytes)
        self.h2d(d, host)
        self.sync()
        return d

    def download(self, dptr, shape, dtype=np.float64):
        out = np.empty(shape, dtype)
        self.d2h(out, dptr)
        self.sync()
        return out

    def sync(self, stream=-1):
        check(lib().bricks_sync(self._h, stream))

    def stream_wait(self, stream, other):
        check(lib().bricks_stream_wait(self._h, stream, other))

    def stream_handle(self, stream):
        p = C.c_void_p()
        check(lib().bricks_stream_handle(self._h, stream, C.byref(p)))
        return p.value

    # timers -----------------------------------------------------------------
    def timer_start(self, slot=0, stream=STREAM_COMPUTE):
        check(lib().bricks_timer_start(self._h, stream, slot))

    def timer_stop(self, slot=0, stream=STREAM_COMPUTE):
        check(lib().bricks_timer_stop(self._h, stream, slot))

    def timer_ms(self, slot=0):
        ms = C.c_float()
        check(lib().bricks_timer_elapsed_ms(self._h, slot, C.byref(ms)))
        return ms.value

    # NCCL -------------------------------------------------------------------
    def comm_init(self, rank, nranks, unique_id, max_ctas=None):
        """max_ctas: cap of the NCCL send/recv kernel's CTAs (bricks_comm_init_config); default from
        BRICKS_NCCL_MAX_CTAS, 0 = NCCL's own choice."""
        buf = C.create_string_buffer(bytes(unique_id), 128)
        if max_ctas is None:
            max_ctas = int(os.environ.get("BRICKS_NCCL_MAX_CTAS", "0"))
        check(lib().bricks_comm_init_config(self._h, rank, nranks, buf, int(max_ctas)))


def unique_id():
    buf = C.create_string_buffer(128)
    check(lib().bricks_comm_unique_id(buf))
    return buf.raw


def launch_count():
    n = C.c_ulonglong()
    check(lib().bricks_launch_count(C.byref(n)))
    return n.value


class Plan(object):
    """Launch plan over boxes of the brick grid (bricks_plan_create).  With `parent`, a plan over
    other boxes of the parent's grid that shares its device tables (bricks_plan_derive; close it
    before the parent).  `borrowed`: a handle owned by a Step."""

    def __init__(self, ctx, grid=None, adj=None, boxes=None, parent=None, borrowed=None):
        self.ctx = ctx
        self._h = C.c_void_p()
        self._owned = borrowed is None
        if borrowed is not None:
            self._h = C.c_void_p(borrowed)
            return
        if parent is not None:
            bx = np.array([list(lo) + list(hi) for lo, hi in boxes], np.uint32)
            check(lib().bricks_plan_derive(parent._h, len(boxes), _np_ptr(bx), C.byref(self._h)))
            self.nbricks = parent.nbricks
        else:
            grid = np.ascontiguousarray(grid, np.uint32)
            adj = np.ascontiguousarray(adj, np.uint32)
            assert grid.ndim == 3 and adj.ndim == 2 and adj.shape[1] == 27
            gd = np.array([grid.shape[2], grid.shape[1], grid.shape[0]], np.uint32)
            if boxes is None:
                boxes = [((0, 0, 0), tuple(int(x) for x in gd))]
            bx = np.array([list(lo) + list(hi) for lo, hi in boxes], np.uint32)
            check(lib().bricks_plan_create(ctx._h, _np_ptr(grid), _np_ptr(gd), _np_ptr(adj), adj.shape[0],
                                           len(boxes), _np_ptr(bx), C.byref(self._h)))
            self.nbricks = adj.shape[0]
        self.ncells = int(sum(np.prod(np.array(hi) - np.array(lo)) for lo, hi in boxes))

    def close(self):
        if self._h and self._owned:
            lib().bricks_plan_destroy(self._h)
        self._h = C.c_void_p()

    def _run(self, fn, stencil, ins, outs, in_step, out_step, coeff, stream):
        if not isinstance(ins, (list, tuple)):
            ins, outs = [ins], [outs]
        n = len(ins)
        ia = (C.c_void_p * n)(*ins)
        oa = (C.c_void_p * n)(*outs)
        cp = None
        if coeff is not None:
            coeff = np.ascontiguousarray(coeff, np.float64)
            cp = _np_ptr(coeff)
        check(fn(self._h, stencil, n, ia, oa, in_step, out_step, cp, stream))

    def device_tables(self):
        """(adj_dev, grid_dev) device pointers held by the plan."""
        a, g = C.c_void_p(), C.c_void_p()
        check(lib().bricks_plan_device_tables(self._h, C.byref(a), C.byref(g)))
        return a.value, g.value

    def stencil(self, stencil, ins, outs, in_step=512, out_step=512, coeff=None, stream=STREAM_COMPUTE):
        self._run(lib().bricks_stencil, stencil, ins, outs, in_step, out_step, coeff, stream)

    def stencil_naive(self, stencil, ins, outs, in_step=512, out_step=512, coeff=None, stream=STREAM_COMPUTE):
        self._run(lib().bricks_stencil_naive, stencil, ins, outs, in_step, out_step, coeff, stream)

    def stencil_f32(self, stencil, ins, outs, in_step=512, out_step=512, coeff=None, stream=STREAM_COMPUTE):
        """fp32 bricks (bElem = float): ins / outs are device pointers to float storages"""
        self._run(lib().bricks_stencil_f32, stencil, ins, outs, in_step, out_step, coeff, stream)

    def stencil_naive_f32(self, stencil, ins, outs, in_step=512, out_step=512, coeff=None, stream=STREAM_COMPUTE):
        self._run(lib().bricks_stencil_naive_f32, stencil, ins, outs, in_step, out_step, coeff, stream)


class Plan4D(object):
    """The 4-D path (weak/main-4d.cpp): grid + 81-slot adjacency of 4x4x4x4 bricks on the device
    (bricks_plan4d_create); sweep() is one brick_stencil call of the reference."""

    def __init__(self, ctx, grid, adj):
        grid = np.ascontiguousarray(grid, np.uint32)
        adj = np.ascontiguousarray(adj, np.uint32)
        assert grid.ndim == 4 and adj.ndim == 2 and adj.shape[1] == 81
        gd = np.array(grid.shape[::-1], np.uint32)
        self.ctx, self.nbricks, self.gdims = ctx, adj.shape[0], tuple(int(x) for x in gd)
        self._h = C.c_void_p()
        check(lib().bricks_plan4d_create(ctx._h, _np_ptr(grid), _np_ptr(gd), _np_ptr(adj), adj.shape[0],
                                         C.byref(self._h)))

    def ncells(self, skip_bricks=0):
        return int(np.prod([max(0, g - 2 * skip_bricks) for g in self.gdims]))

    def sweep(self, in_dev, out_dev, skip_bricks=0, in_step=256, out_step=256, stream=STREAM_COMPUTE):
        check(lib().bricks_stencil4d(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev), in_step, out_step,
                                     skip_bricks, stream))

    def sweep_naive(self, in_dev, out_dev, skip_bricks=0, in_step=256, out_step=256, stream=STREAM_COMPUTE):
        check(lib().bricks_stencil4d_naive(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev), in_step, out_step,
                                           skip_bricks, stream))

    def close(self):
        if self._h:
            lib().bricks_plan4d_destroy(self._h)
            self._h = C.c_void_p()


class Exchange(object):
    """Halo exchange plan (bricks_exchange_create).  Offsets and sizes in bytes."""

    def __init__(self, ctx, rank, recv_off, send_off, nbytes, recv_peer, send_peer):
        a = [np.ascontiguousarray(x, np.uint64) for x in (recv_off, send_off, nbytes)]
        p = [np.ascontiguousarray(x, np.int32) for x in (recv_peer, send_peer)]
        self._h = C.c_void_p()
        self.ctx = ctx
        check(lib().bricks_exchange_create(ctx._h, rank, len(a[0]), _np_ptr(a[0]), _np_ptr(a[1]), _np_ptr(a[2]),
                                           _np_ptr(p[0]), _np_ptr(p[1]), C.byref(self._h)))

    def run(self, storage_dev, stream=STREAM_COMM):
        check(lib().bricks_exchange_run(self._h, C.c_void_p(storage_dev), stream))

    def close(self):
        if self._h:
            lib().bricks_exchange_destroy(self._h)
            self._h = C.c_void_p()


class ShiftExchange(object):
    """MultiStageExchangeView (brick-mpi.h:126-176): three Exchange plans, one per axis, run in order
    on one stream.  Build from Decomp.shift_plan()."""

    def __init__(self, ctx, rank, stages):
        self.stages = [Exchange(ctx, rank, *st) for st in stages]
        self.npieces = [len(st[0]) for st in stages]

    def run(self, storage_dev, stream=STREAM_COMM):
        for ex in self.stages:
            ex.run(storage_dev, stream)

    def close(self):
        for ex in self.stages:
            ex.close()
        self.stages = []


PLAN_FULL, PLAN_INNER, PLAN_SHELL, PLAN_OWNED, PLAN_OSKIN = range(5)
BASE_ABS, BASE_FIELD, BASE_ARENA = 0, 1, 2
STEP_OVERLAP, STEP_EARLY, STEP_EARLY_UNSPLIT = 1, 2, 4


def step_boxes(gdims, which, ghost_bricks=1):
    """the boxes of plan `which` of the library's sweep schedule (bricks_step_boxes; host only)"""
    gd = np.array(list(gdims), np.uint32)
    bx = np.zeros(36, np.uint32)
    n = C.c_int()
    check(lib().bricks_step_boxes(_np_ptr(gd), ghost_bricks, which, _np_ptr(bx), C.byref(n)))
    return [(tuple(int(x) for x in bx[6 * b:6 * b + 3]), tuple(int(x) for x in bx[6 * b + 3:6 * b + 6]))
            for b in range(n.value)]


class Step(object):
    """The sweep schedule of one rank (bricks_step_*): five box plans over one device copy of
    grid / adj plus the communication program (copy tables and exchange plans)."""

    def __init__(self, ctx, grid=None, adj=None, ghost_bricks=1, handle=None):
        self.ctx = ctx
        self._h = C.c_void_p()
        if handle is not None:  # made by Decomp.make_step
            self._h = C.c_void_p(handle)
            return
        grid = np.ascontiguousarray(grid, np.uint32)
        adj = np.ascontiguousarray(adj, np.uint32)
        gd = np.array([grid.shape[2], grid.shape[1], grid.shape[0]], np.uint32)
        check(lib().bricks_step_create(ctx._h, _np_ptr(grid), _np_ptr(gd), _np_ptr(adj), adj.shape[0],
                                       ghost_bricks, C.byref(self._h)))

    def plan(self, which):
        """borrowed Plan handle, or None when the subdomain has no inner box"""
        h = C.c_void_p()
        if lib().bricks_step_plan(self._h, which, C.byref(h)) != 0:
            return None
        return Plan(self.ctx, borrowed=h.value)

    def arena(self, nbytes):
        p = C.c_void_p()
        check(lib().bricks_step_arena(self._h, nbytes, C.byref(p)))
        return p.value

    def add_copy(self, links, src_base=BASE_ABS, dst_base=BASE_ABS):
        """links: list of (from, to, nbytes); from / to are pointers or offsets from the base"""
        if not links:
            return
        arr = (RegionLink * len(links))(*[RegionLink(s, d, n) for s, d, n in links])
        check(lib().bricks_step_add_copy(self._h, arr, len(links), src_base, dst_base))

    def add_exchange(self, ex, base=BASE_FIELD):
        check(lib().bricks_step_add_exchange(self._h, ex._h, base, 1))
        ex._h = C.c_void_p()  # the step owns it now

    @staticmethod
    def _ptrs(x):
        if not isinstance(x, (list, tuple)):
            x = [x]
        return (C.c_void_p * len(x))(*x), len(x)

    def run(self, stencil, st_iter, a, b, coeff=None, flags=STEP_OVERLAP):
        pa, n = self._ptrs(a)
        pb, _ = self._ptrs(b)
        cp = None
        if coeff is not None:
            coeff = np.ascontiguousarray(coeff, np.float64)
            cp = _np_ptr(coeff)
        check(lib().bricks_step_run(self._h, stencil, st_iter, n, pa, pb, cp, flags))

    def exchange(self, field0):
        check(lib().bricks_step_exchange(self._h, C.c_void_p(field0)))

    def sweep(self, which, stencil, ins, outs, coeff=None, stream=STREAM_COMPUTE):
        pa, n = self._ptrs(ins)
        pb, _ = self._ptrs(outs)
        cp = None
        if coeff is not None:
            coeff = np.ascontiguousarray(coeff, np.float64)
            cp = _np_ptr(coeff)
        check(lib().bricks_step_sweep(self._h, which, stencil, n, pa, pb, cp, stream))

    def reset(self):
        check(lib().bricks_step_reset(self._h))

    def p2p_export(self, field0, rank):
        """256-byte descriptor of the storage this rank exchanges (bricks_step_p2p_export)"""
        buf = C.create_string_buffer(256)
        check(lib().bricks_step_p2p_export(self._h, C.c_void_p(field0), rank, buf))
        return buf.raw

    def p2p_connect(self, descs):
        """descs: the descriptors of ALL ranks, in rank order (bricks_step_p2p_connect)"""
        blob = b"".join(descs)
        assert len(blob) == 256 * len(descs)
        check(lib().bricks_step_p2p_connect(self._h, len(descs), C.create_string_buffer(blob, len(blob))))

    def p2p_disable(self):
        check(lib().bricks_step_p2p_disable(self._h))

    def trace(self, enable=True):
        check(lib().bricks_step_trace(self._h, 1 if enable else 0))

    def trace_read(self):
        """timeline of the last traced step: list of (name, start_ms, end_ms), synchronises"""
        n = C.c_int()
        check(lib().bricks_step_trace_read(self._h, 0, C.byref(n), None, None, None))
        if n.value == 0:
            return []
        codes = np.zeros(n.value, np.int32)
        t0, t1 = np.zeros(n.value, np.float32), np.zeros(n.value, np.float32)
        check(lib().bricks_step_trace_read(self._h, n.value, C.byref(n), _np_ptr(codes), _np_ptr(t0), _np_ptr(t1)))
        names = {PLAN_FULL: "full", PLAN_INNER: "inner", PLAN_SHELL: "shell", PLAN_OWNED: "owned", PLAN_OSKIN: "oskin"}
        out = []
        for c, a, b in zip(codes.tolist(), t0.tolist(), t1.tolist()):
            if c >= 3000:
                name = "comm[%d] exchange" % (c - 3000)
            elif c >= 2000:
                name = "comm[%d] copy table" % (c - 2000)
            else:
                name = "sweep %d %s" % ((c - 1000) // 10 + 1, names[(c - 1000) % 10])
            out.append((name, a, b))
        return out

    def close(self):
        if self._h:
            lib().bricks_step_destroy(self._h)
            self._h = C.c_void_p()


COPY_STATIC_MAX_LINKS = 2048  # BRICKS_COPY_STATIC_MAX_LINKS


def copy_regions_static(ctx, links, stream=STREAM_COMPUTE):
    """bricks_copy_regions_static: the same copies without any allocation (the table is staged in a
    static device array of the library).  links: list of (src_dev, dst_dev, nbytes)."""
    arr = (RegionLink * max(len(links), 1))(*[RegionLink(s, d, n) for s, d, n in links])
    check(lib().bricks_copy_regions_static(ctx._h, arr, len(links), stream))


def copy_regions(ctx, links, stream=STREAM_COMPUTE):
    """links: list of (src_dev, dst_dev, nbytes).  Uploads the table then runs the copy kernel."""
    arr = (RegionLink * len(links))(*[RegionLink(s, d, n) for s, d, n in links])
    dev = C.c_void_p()
    check(lib().bricks_links_upload(ctx._h, arr, len(links), C.byref(dev)))
    check(lib().bricks_copy_regions(ctx._h, dev, len(links), stream))
    ctx.sync(stream)  # the table is read by the kernel; free it once the copies are done
    check(lib().bricks_links_free(ctx._h, dev))


# probes ---------------------------------------------------------------------
def probe_dfma(ctx, iters=20000):
    t, ms = C.c_double(), C.c_float()
    check(lib().bricks_probe_dfma(ctx._h, iters, C.byref(t), C.byref(ms)))
    return t.value


def probe_copy(ctx, nbytes, reps=10):
    g = C.c_double()
    check(lib().bricks_probe_copy(ctx._h, nbytes, reps, C.byref(g)))
    return g.value


def probe_read(ctx, nbytes, reps=10):
    g = C.c_double()
    check(lib().bricks_probe_read(ctx._h, nbytes, reps, C.byref(g)))
    return g.value


def probe_mio(ctx, mode):
    a, b = C.c_double(), C.c_double()
    check(lib().bricks_probe_mio(ctx._h, mode, C.byref(a), C.byref(b)))
    return a.value, b.value


def probe_tma_issue(ctx, field_dev, nbricks, mode, nops=240, step=512):
    a, b = C.c_double(), C.c_double()
    check(lib().bricks_probe_tma_issue(ctx._h, C.c_void_p(field_dev), nbricks, step, mode, nops, C.byref(a), C.byref(b)))
    return a.value, b.value


def probe_tma_image(ctx, field_dev, nbricks, step, b):
    out = np.zeros(512, np.float64)
    check(lib().bricks_probe_tma_image(ctx._h, C.c_void_p(field_dev), nbricks, step, b, _np_ptr(out)))
    return out


# ---------------------------------------------------------------------------
# host templates through include/bricks_b200_host.h (no device involved)
REGION_DT = np.dtype([("neighbor", np.uint64), ("skin_st", np.uint32), ("skin_ed", np.uint32),
                      ("pos", np.uint32), ("len", np.uint32), ("first_pad", np.uint32),
                      ("last_pad", np.uint32)], align=True)
SKIN_GOOD, SKIN_NORMAL, SKIN_BAD = 0, 1, 2

_SIGS.update({
    "bricks_decomp_create": (C.c_int, [_vp, C.c_uint, C.c_int, _vpp]),
    "bricks_decomp_destroy": (C.c_int, [_vp]),
    "bricks_decomp_sizes": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp_tables": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp_rank_map": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp_exchange_plan": (C.c_int, [_vp, C.c_size_t, _vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp_shift_plan": (C.c_int, [_vp, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_int), _vp, _vp, _vp, _vp, _vp]),
    "bricks_init_grid3": (C.c_int, [_vp, _vp, _vp]),
    "bricks_copy_brick3": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp, C.c_uint]),
    "bricks_compare_brick3": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp, C.c_uint, _vp, _vp]),
    "bricks_cart_dims": (C.c_int, [C.c_int, C.c_int, _vp]),
    "bricks_cart_rank": (C.c_int, [C.c_int, _vp, _vp, _vp]),
    "bricks_cart_coords": (C.c_int, [C.c_int, _vp, C.c_int, _vp]),
    "bricks_zmort3": (C.c_ulong, [C.c_ulong, C.c_ulong, C.c_ulong]),
    "bricks_zmort3_coord": (C.c_ulong, [C.c_ulong, C.c_ulong]),
    "bricks_zmort3_set": (C.c_ulong, [C.c_ulong, C.c_ulong, C.c_ulong]),
    "bricks_uniform_fill": (C.c_int, [_vp, C.c_size_t, C.c_uint64, C.c_uint64]),
    "bricks_decomp4d_create": (C.c_int, [_vp, C.c_uint, C.c_int, _vpp]),
    "bricks_decomp4d_destroy": (C.c_int, [_vp]),
    "bricks_decomp4d_sizes": (C.c_int, [_vp, C.POINTER(C.c_uint), _vp, _vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "bricks_decomp4d_tables": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp4d_rank_map": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "bricks_decomp4d_exchange_plan": (C.c_int, [_vp, C.c_size_t, _vp, _vp, _vp, _vp, _vp]),
    "bricks_copy_brick4": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp, C.c_uint]),
})


class Decomp(object):
    """BrickDecomp<3,8,8,8>(dims, depth).initialize(skin3d_*) -- the product's own C++
    template (include/bricks/brick-decomp.h), reached through bricks_decomp_*."""

    def __init__(self, dims, depth=8, skin=SKIN_GOOD):
        L = lib()
        d = np.asarray(dims, np.uint32)
        self._h = C.c_void_p()
        check(L.bricks_decomp_create(_np_ptr(d), depth, skin, C.byref(self._h)))
        self.dims = tuple(int(x) for x in dims)
        self.depth = depth
        nb, nr, ns = C.c_uint(), C.c_int(), C.c_int()
        self.sep_pos = np.zeros(3, np.uint32)
        gd = np.zeros(3, np.uint32)
        check(L.bricks_decomp_sizes(self._h, C.byref(nb), _np_ptr(self.sep_pos), _np_ptr(gd), C.byref(nr), C.byref(ns)))
        self.nbricks, self.nregions = nb.value, nr.value
        self.gdims = tuple(int(x) for x in gd)  # contiguous axis first
        grid = np.zeros(int(gd[0]) * int(gd[1]) * int(gd[2]), np.uint32)
        self.adj = np.zeros((self.nbricks, 27), np.uint32)
        self.ghost = np.zeros(nr.value, REGION_DT)
        self.skin = np.zeros(nr.value, REGION_DT)
        self.skin_size = np.zeros(ns.value, np.int64)
        assert REGION_DT.itemsize == 32
        check(L.bricks_decomp_tables(self._h, _np_ptr(grid), _np_ptr(self.adj), _np_ptr(self.ghost),
                                     _np_ptr(self.skin), _np_ptr(self.skin_size)))
        self.grid = grid.reshape(int(gd[2]), int(gd[1]), int(gd[0]))

    def rank_map(self, procdims, coo):
        """populate(): {direction set: rank} on the periodic process grid `procdims`."""
        keys, ranks = np.zeros(27, np.uint64), np.zeros(27, np.int32)
        pd, cc = np.asarray(procdims, np.int32), np.asarray(coo, np.int32)
        check(lib().bricks_decomp_rank_map(self._h, _np_ptr(pd), _np_ptr(cc), _np_ptr(keys), _np_ptr(ranks)))
        return {int(k): int(r) for k, r in zip(keys, ranks)}

    def exchange_plan(self, step=512):
        """(recv_off, send_off, nbytes, recv_peer, send_peer) of BrickDecomp::exchange for the
        rank map set by the last rank_map() call."""
        n = self.nregions
        ro, so, nb = (np.zeros(n, np.uint64) for _ in range(3))
        rp, sp = np.zeros(n, np.int32), np.zeros(n, np.int32)
        check(lib().bricks_decomp_exchange_plan(self._h, step, _np_ptr(ro), _np_ptr(so), _np_ptr(nb),
                                                _np_ptr(rp), _np_ptr(sp)))
        return ro, so, nb, rp, sp

    def make_step(self, ctx, rank, nranks, kind="regions", step=512):
        """BrickDecomp::makeStep: the library's sweep schedule for this decomposition with the
        communication program of exchange `kind` ("regions" | "packed" | "shift"); call rank_map()
        first.  Returns (Step, description)."""
        k = {"regions": 0, "packed": 1, "shift": 2}[kind]
        h = C.c_void_p()
        what = C.create_string_buffer(200)
        check(lib().bricks_decomp_step_create(self._h, ctx._h, rank, nranks, k, step, C.byref(h), what, 200))
        return Step(ctx, handle=h.value), what.value.decode()

    def shift_plan(self, step=512):
        """The SHIFT exchange (multiStageExchangeView, brick-mpi.h:587-681) for the rank map set by the
        last rank_map() call: three (recv_off, send_off, nbytes, recv_peer, send_peer) tuples, one
        per axis, each the arguments of an Exchange; stages must run in order."""
        out = []
        for stage in range(3):
            n = C.c_int()
            check(lib().bricks_decomp_shift_plan(self._h, step, stage, 0, C.byref(n), None, None, None, None, None))
            ro, so, nb = (np.zeros(n.value, np.uint64) for _ in range(3))
            rp, sp = np.zeros(n.value, np.int32), np.zeros(n.value, np.int32)
            check(lib().bricks_decomp_shift_plan(self._h, step, stage, n.value, C.byref(n), _np_ptr(ro), _np_ptr(so),
                                                 _np_ptr(nb), _np_ptr(rp), _np_ptr(sp)))
            out.append((ro, so, nb, rp, sp))
        return out

    def close(self):
        if self._h:
            lib().bricks_decomp_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Decomp4D(object):
    """BrickDecomp<4, 4,4,4,4>(dims, depth).initialize(allneighbors minus the centre) of
    weak/main-4d.cpp:98-106 -- the product's own C++ template through bricks_decomp4d_*."""

    def __init__(self, dims, depth=8, page_align=False):
        L = lib()
        d = np.asarray(dims, np.uint32)
        assert len(d) == 4
        self._h = C.c_void_p()
        check(L.bricks_decomp4d_create(_np_ptr(d), depth, 1 if page_align else 0, C.byref(self._h)))
        self.dims, self.depth = tuple(int(x) for x in dims), depth
        nb, nr, ns = C.c_uint(), C.c_int(), C.c_int()
        self.sep_pos = np.zeros(3, np.uint32)
        gd = np.zeros(4, np.uint32)
        check(L.bricks_decomp4d_sizes(self._h, C.byref(nb), _np_ptr(self.sep_pos), _np_ptr(gd), C.byref(nr),
                                      C.byref(ns)))
        self.nbricks, self.nregions = nb.value, nr.value
        self.gdims = tuple(int(x) for x in gd)  # contiguous axis first
        self.grid = np.zeros(self.gdims[::-1], np.uint32)
        self.adj = np.zeros((self.nbricks, 81), np.uint32)
        self.ghost = np.zeros(nr.value, REGION_DT)
        self.skin = np.zeros(nr.value, REGION_DT)
        self.skin_size = np.zeros(ns.value, np.int64)
        check(L.bricks_decomp4d_tables(self._h, _np_ptr(self.grid), _np_ptr(self.adj), _np_ptr(self.ghost),
                                       _np_ptr(self.skin), _np_ptr(self.skin_size)))

    def rank_map(self, procdims, coo):
        keys, ranks = np.zeros(81, np.uint64), np.zeros(81, np.int32)
        pd, cc = np.asarray(procdims, np.int32), np.asarray(coo, np.int32)
        check(lib().bricks_decomp4d_rank_map(self._h, _np_ptr(pd), _np_ptr(cc), _np_ptr(keys), _np_ptr(ranks)))
        return {int(k): int(r) for k, r in zip(keys, ranks)}

    def exchange_plan(self, step=256):
        """(recv_off, send_off, nbytes, recv_peer, send_peer): the arguments of an Exchange, for the rank
        map set by the last rank_map() call"""
        n = self.nregions
        ro, so, nb = (np.zeros(n, np.uint64) for _ in range(3))
        rp, sp = np.zeros(n, np.int32), np.zeros(n, np.int32)
        check(lib().bricks_decomp4d_exchange_plan(self._h, step, _np_ptr(ro), _np_ptr(so), _np_ptr(nb),
                                                  _np_ptr(rp), _np_ptr(sp)))
        return ro, so, nb, rp, sp

    def close(self):
        if self._h:
            lib().bricks_decomp4d_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def copy_brick4(to_brick, dimlist, padding, ghost, arr, grid, bdat, adj, step=256):
    """copyToBrick<4> / copyFromBrick<4> between a padded 4-D array and row-major 4x4x4x4 bricks."""
    d, p, g = (np.asarray(x, np.int64) for x in (dimlist, padding, ghost))
    grid = np.ascontiguousarray(grid, np.uint32)
    adj = np.ascontiguousarray(adj, np.uint32)
    assert arr.flags["C_CONTIGUOUS"] and bdat.flags["C_CONTIGUOUS"]
    check(lib().bricks_copy_brick4(1 if to_brick else 0, _np_ptr(d), _np_ptr(p), _np_ptr(g), _np_ptr(arr),
                                   _np_ptr(grid), _np_ptr(bdat), step, _np_ptr(adj), adj.shape[0]))


def init_grid(dimlist):
    """init_grid<3>: (grid[k][j][i], adj) with identity numbering (bricksetup.h:73-90)."""
    d = np.asarray(dimlist, np.int64)
    n = int(d[0] * d[1] * d[2])
    grid, adj = np.zeros(n, np.uint32), np.zeros((n, 27), np.uint32)
    check(lib().bricks_init_grid3(_np_ptr(d), _np_ptr(grid), _np_ptr(adj)))
    return grid.reshape(int(d[2]), int(d[1]), int(d[0])), adj


def copy_brick(to_brick, dimlist, padding, ghost, arr, grid, bdat, adj, step=512):
    """copyToBrick<3> / copyFromBrick<3> between a padded array and row-major bricks."""
    d, p, g = (np.asarray(x, np.int64) for x in (dimlist, padding, ghost))
    grid = np.ascontiguousarray(grid, np.uint32)
    adj = np.ascontiguousarray(adj, np.uint32)
    assert arr.flags["C_CONTIGUOUS"] and bdat.flags["C_CONTIGUOUS"]
    check(lib().bricks_copy_brick3(1 if to_brick else 0, _np_ptr(d), _np_ptr(p), _np_ptr(g), _np_ptr(arr),
                                   _np_ptr(grid), _np_ptr(bdat), step, _np_ptr(adj), adj.shape[0]))


def copy_array_dev(ctx, to_brick, dimlist, padding, ghost, arr_dev, grid_dev, bricks_dev, step=512,
                   stream=STREAM_COMPUTE):
    """copyToBrick<3> / copyFromBrick<3> on the device (bricks_copy_array): array and bricks are device
    pointers, grid_dev the device brick-index grid (Plan.device_tables()[1])."""
    d, p, g = (np.ascontiguousarray(x, np.int64) for x in (dimlist, padding, ghost))
    check(lib().bricks_copy_array(ctx._h, 1 if to_brick else 0, _np_ptr(d), _np_ptr(p), _np_ptr(g), arr_dev,
                                  grid_dev, bricks_dev, step, stream))


def array_stencil7(ctx, in_dev, out_dev, ext, lo, hi, coeff, stream=STREAM_COMPUTE):
    """Array-layout 7-point baseline on the device (d3pt7_arr of stencils/3axis.cu:39-49)."""
    e, l, h = (np.ascontiguousarray(x, np.int64) for x in (ext, lo, hi))
    c = np.ascontiguousarray(coeff, np.float64)
    check(lib().bricks_array_stencil7(ctx._h, in_dev, out_dev, _np_ptr(e), _np_ptr(l), _np_ptr(h), _np_ptr(c), stream))


def compare_brick(dimlist, padding, ghost, arr, grid, bdat, adj, step=512):
    """(compareBrick verdict at BRICK_TOLERANCE, strict max relative error)."""
    d, p, g = (np.asarray(x, np.int64) for x in (dimlist, padding, ghost))
    grid = np.ascontiguousarray(grid, np.uint32)
    adj = np.ascontiguousarray(adj, np.uint32)
    m, r = C.c_int(), C.c_double()
    check(lib().bricks_compare_brick3(_np_ptr(d), _np_ptr(p), _np_ptr(g), _np_ptr(arr), _np_ptr(grid),
                                      _np_ptr(bdat), step, _np_ptr(adj), adj.shape[0], C.byref(m), C.byref(r)))
    return bool(m.value), r.value


def cart_dims(nranks, ndims=3):
    out = np.zeros(ndims, np.int32)
    check(lib().bricks_cart_dims(nranks, ndims, _np_ptr(out)))
    return [int(x) for x in out]


def cart_coords(rank, dims):
    d = np.asarray(dims, np.int32)
    out = np.zeros(len(dims), np.int32)
    check(lib().bricks_cart_coords(len(dims), _np_ptr(d), rank, _np_ptr(out)))
    return [int(x) for x in out]


def cart_rank(coo, dims):
    d, c = np.asarray(dims, np.int32), np.asarray(coo, np.int32)
    r = C.c_int()
    check(lib().bricks_cart_rank(len(dims), _np_ptr(d), _np_ptr(c), C.byref(r)))
    return r.value


def uniform(n, seed=0, start=0):
    """randomArray(list, seed) of include/bricks/multiarray.h (== bricklib_b200.synth.uniform)."""
    out = np.empty(n, np.float64)
    check(lib().bricks_uniform_fill(_np_ptr(out), n, seed, start))
    return out


# ---------------------------------------------------------------------------
# strong scaling (include/bricks/strong-plan.h through bricks_strong_*) and reusable link tables
class _StrongLink(C.Structure):
    _fields_ = [("src_sub", C.c_uint), ("dst_sub", C.c_uint), ("src_off", C.c_size_t), ("dst_off", C.c_size_t),
                ("len", C.c_size_t)]


class _StrongPiece(C.Structure):
    _fields_ = [("sub", C.c_uint), ("region", C.c_uint), ("off", C.c_size_t), ("len", C.c_size_t),
                ("buf_off", C.c_size_t), ("sender_id", C.c_ulong)]


_SIGS.update({
    "bricks_exchange_create_v": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vpp]),
    "bricks_strong_plan_create": (C.c_int, [C.c_uint, C.c_uint, C.c_int, C.c_int, _vp, C.c_size_t, _vpp]),
    "bricks_strong_plan_destroy": (C.c_int, [_vp]),
    "bricks_strong_plan_sizes": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "bricks_strong_plan_links": (C.c_int, [_vp, _vp]),
    "bricks_strong_plan_peer": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp]),
    "bricks_strong_plan_pieces": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "bricks_strong_owner": (C.c_int, [_vp, C.c_ulong, _vp]),
    "bricks_strong_neighbor": (C.c_int, [_vp, C.c_ulong, C.c_uint64, _vp]),
    "bricks_strong_coords": (C.c_int, [_vp, C.c_ulong, _vp]),
})


class StrongPlan(object):
    """Ownership and exchange lists of the two-level decomposition (strong/args.cpp:36-56,104-113,
    strong/main.cu:188-395), from the product's C++ planner."""

    def __init__(self, dom, sdom, nranks, rank, dec, step=512):
        L = lib()
        self._h = C.c_void_p()
        check(L.bricks_strong_plan_create(dom, sdom, nranks, rank, dec._h, step, C.byref(self._h)))
        lo, hi, nl = C.c_ulong(), C.c_ulong(), C.c_size_t()
        ns, nr = C.c_int(), C.c_int()
        check(L.bricks_strong_plan_sizes(self._h, C.byref(lo), C.byref(hi), C.byref(nl), C.byref(ns), C.byref(nr)))
        self.first_id, self.last_id = lo.value, hi.value
        self.nsub = hi.value - lo.value
        arr = (_StrongLink * max(nl.value, 1))()
        check(L.bricks_strong_plan_links(self._h, arr))
        self.links = [(l.src_sub, l.dst_sub, l.src_off, l.dst_off, l.len) for l in arr[:nl.value]]
        self.send, self.recv = [], []
        for d, out, n in ((0, self.send, ns.value), (1, self.recv, nr.value)):
            for i in range(n):
                rk, by, npc = C.c_int(), C.c_size_t(), C.c_int()
                check(L.bricks_strong_plan_peer(self._h, d, i, C.byref(rk), C.byref(by), C.byref(npc)))
                pcs = (_StrongPiece * max(npc.value, 1))()
                check(L.bricks_strong_plan_pieces(self._h, d, i, pcs))
                out.append({"rank": rk.value, "bytes": by.value,
                            "pieces": [(p.sub, p.region, p.off, p.len, p.buf_off, p.sender_id) for p in pcs[:npc.value]]})

    def owner(self, sid):
        r = C.c_int()
        check(lib().bricks_strong_owner(self._h, sid, C.byref(r)))
        return r.value

    def neighbor(self, sid, direction_set):
        n = C.c_ulong()
        check(lib().bricks_strong_neighbor(self._h, sid, direction_set, C.byref(n)))
        return n.value

    def coords(self, sid):
        c = (C.c_uint * 3)()
        check(lib().bricks_strong_coords(self._h, sid, c))
        return (c[0], c[1], c[2])

    def close(self):
        if self._h:
            lib().bricks_strong_plan_destroy(self._h)
            self._h = C.c_void_p()


class LinkTable(object):
    """A device-resident RegionLink list (bricks_links_upload) that can be run many times."""

    def __init__(self, ctx, links):
        self.ctx, self.n = ctx, len(links)
        self._dev = C.c_void_p()
        if self.n:
            arr = (RegionLink * self.n)(*[RegionLink(s, d, n) for s, d, n in links])
            check(lib().bricks_links_upload(ctx._h, arr, self.n, C.byref(self._dev)))
        self.bytes = sum(n for _, _, n in links)

    def run(self, stream=STREAM_COMPUTE):
        if self.n:
            check(lib().bricks_copy_regions(self.ctx._h, self._dev, self.n, stream))

    def close(self):
        if self._dev:
            lib().bricks_links_free(self.ctx._h, self._dev)
            self._dev = C.c_void_p()


class ExchangeV(Exchange):
    """bricks_exchange_create_v: independent receive / send sizes per entry."""

    def __init__(self, ctx, rank, recv_off, recv_bytes, recv_peer, send_off, send_bytes, send_peer):
        a = [np.ascontiguousarray(x, np.uint64) for x in (recv_off, recv_bytes, send_off, send_bytes)]
        p = [np.ascontiguousarray(x, np.int32) for x in (recv_peer, send_peer)]
        self._h = C.c_void_p()
        self.ctx = ctx
        check(lib().bricks_exchange_create_v(ctx._h, rank, len(a[0]), _np_ptr(a[0]), _np_ptr(a[1]), _np_ptr(p[0]),
                                             _np_ptr(a[2]), _np_ptr(a[3]), _np_ptr(p[1]), C.byref(self._h)))
