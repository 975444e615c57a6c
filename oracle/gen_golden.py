"""oracle/gen_golden.py -- TEST INFRASTRUCTURE, not product code.

Generates tests/golden/* from the UNMODIFIED reference compiled into oracle/_ref by
oracle/Makefile (BrickDecomp::initialize / populate of include/brick-mpi.h, the codegen
brick kernels of stencils/*.py and the ST_CPU array forms of stencils/fake.h).  Run it in
the build container, where /root/reference exists:

    make -C oracle all && python oracle/gen_golden.py && python oracle/gen_golden.py cond
    python oracle/gen_golden.py d4        # the 4-D path

The fixtures pin the oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.
Inputs are not stored: they are a pure function of the element index
(bricklib_b200/synth.py), so every machine regenerates them bit for bit.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from bricklib_b200 import synth  # noqa: E402
from oracle import refload as R  # noqa: E402

GOLDEN = os.path.join(HERE, "..", "tests", "golden")
STENCILS = (0, 7, 13, 25, 125)
DECOMP_DIMS = [(16, 16, 16), (32, 16, 24), (64, 64, 64), (128, 128, 128), (512, 512, 512)]
RANK_CASES = [((1, 1, 1), (0, 0, 0)), ((2, 1, 1), (1, 0, 0)), ((2, 2, 1), (0, 1, 0)),
              ((2, 2, 2), (0, 1, 1)), ((2, 2, 2), (1, 0, 1)), ((3, 2, 1), (2, 1, 0))]
STENCIL_DIMS = [(16, 16, 16), (16, 24, 32)]


def fnv1a64(values):
    """FNV-1a over a stream of integers, one value per step."""
    h = 0xCBF29CE484222325
    for v in np.asarray(values).reshape(-1).tolist():
        h ^= int(v)
        h = (h * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def decomp_record(dims):
    ref = R.RefDecomp(dims, 8)
    rec = {"dims": list(dims), "gz": 8, "nbricks": int(ref.nbricks),
           "sep_pos": [int(x) for x in ref.sep_pos],
           "skin_size": [int(x) for x in ref.skin_size],
           "adj_fnv": fnv1a64(ref.adj), "grid_fnv": fnv1a64(ref.grid)}
    for name in ("ghost", "skin"):
        rec[name] = {k: [int(x) for x in v] for k, v in ref.regions[name].items()}
    rec["rank_map"] = []
    for procdims, coo in RANK_CASES:
        rm = ref.rank_map(procdims, coo)
        rec["rank_map"].append({"procdims": list(procdims), "coo": list(coo),
                                "map": {"%x" % k: v for k, v in sorted(rm.items())}})
    ref.close()
    return rec


def stencil_fixture(dims):
    """Reference outputs on one decomposed subdomain: codegen brick kernel over every grid
    brick (stored for the interior bricks only, the region the reference itself compares),
    and the ST_CPU array form over the interior of the matching padded array."""
    ref = R.RefDecomp(dims, 8)
    nb = ref.nbricks
    inp = synth.uniform(nb * 512, seed=11)
    coeff = synth.coefficients()
    blist = ref.grid.reshape(-1)
    t = ref.grid.shape  # (tk, tj, ti) brick grid incl. ghost
    interior = ref.grid[1:-1, 1:-1, 1:-1].reshape(-1)
    out = {"dims": np.array(dims), "interior_bricks": interior.astype(np.uint32)}
    # padded array twin of the same field: element (k,j,i) of the (dims+16)^3 region
    arr = np.zeros((t[0] * 8, t[1] * 8, t[2] * 8))
    field = inp.reshape(nb, 8, 8, 8)
    for bk in range(t[0]):
        for bj in range(t[1]):
            for bi in range(t[2]):
                arr[bk * 8:bk * 8 + 8, bj * 8:bj * 8 + 8, bi * 8:bi * 8 + 8] = field[ref.grid[bk, bj, bi]]
    for w in STENCILS:
        cf = coeff if w == 0 else None
        res = np.zeros(nb * 512)
        R.brick_sweep_rm(w, ref.adj, blist, inp, res, 512, cf)
        out["brick_%d" % w] = res.reshape(nb, 512)[interior].copy()
        aout = np.zeros_like(arr)
        box = (8, arr.shape[0] - 8, 8, arr.shape[1] - 8, 8, arr.shape[2] - 8)
        R.array_sweep(w, arr, aout, box, cf)
        out["array_%d" % w] = aout[8:-8, 8:-8, 8:-8].copy()
    ref.close()
    return out


def cond_inputs(nelem):
    """Field and coefficients with both signs, so that max(x, 0) and |sum| of cond.py matter."""
    return synth.uniform(nelem, seed=13) - 0.5, synth.coefficients()[:7] - 0.5


def cond_fixture(dims):
    """cond.py (d3cond, stencils/3axis.cpp:94-163) through the reference's codegen brick kernel and
    its array form, on a decomposed subdomain; same layout as stencil_fixture."""
    ref = R.RefDecomp(dims, 8)
    nb = ref.nbricks
    inp, coeff = cond_inputs(nb * 512)
    t = ref.grid.shape
    interior = ref.grid[1:-1, 1:-1, 1:-1].reshape(-1)
    arr = np.zeros((t[0] * 8, t[1] * 8, t[2] * 8))
    field = inp.reshape(nb, 8, 8, 8)
    for bk in range(t[0]):
        for bj in range(t[1]):
            for bi in range(t[2]):
                arr[bk * 8:bk * 8 + 8, bj * 8:bj * 8 + 8, bi * 8:bi * 8 + 8] = field[ref.grid[bk, bj, bi]]
    res = np.zeros(nb * 512)
    R.brick_sweep_rm(1, ref.adj, ref.grid.reshape(-1), inp, res, 512, coeff)
    aout = np.zeros_like(arr)
    R.array_sweep(1, arr, aout, (8, arr.shape[0] - 8, 8, arr.shape[1] - 8, 8, arr.shape[2] - 8), coeff)
    out = {"dims": np.array(dims), "interior_bricks": interior.astype(np.uint32),
           "brick_1": res.reshape(nb, 512)[interior].copy(), "array_1": aout[8:-8, 8:-8, 8:-8].copy()}
    ref.close()
    return out


def d3pt27_fixture():
    """Runs the reference's own d3pt27() (stencils/3axis.cpp:165-247, N = 256, its own random field)
    through oracle/refshim/ref_3axis.cpp (`make -C oracle golden27`) and cuts one 16^3 block of its
    output together with the 18^3 block of its input that the block depends on.  d3pt27() throws
    when its brick form and its array form disagree, so the block stands for both."""
    import ctypes as C
    import subprocess
    from bricklib_b200 import synth
    subprocess.check_call(["make", "-C", HERE, "golden27"], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HERE, "_ref", "libref3axis.so"))
    n, stride, first = C.c_long(), C.c_long(), C.c_long()
    lib.ref3axis_geometry(C.byref(n), C.byref(stride), C.byref(first))
    coeff = np.ascontiguousarray(synth.coefficients()[:27] - 0.5)  # both signs
    pin, pout = C.POINTER(C.c_double)(), C.POINTER(C.c_double)()
    rc = lib.ref3axis_d3pt27(coeff.ctypes.data_as(C.POINTER(C.c_double)), C.byref(pin), C.byref(pout))
    if rc:
        raise RuntimeError("reference d3pt27 failed: %d" % rc)
    s, f = stride.value, first.value
    arr = np.ctypeslib.as_array(pin, shape=(s, s, s))
    out = np.ctypeslib.as_array(pout, shape=(s, s, s))
    assert out[0, 0, 0] == 0.0 and out[f, f, f] != 0.0  # zeroArray outside the computed box
    k0, j0, i0 = f + 40, f + 8, f + 100  # a block off every symmetry plane, inside the domain
    fx = {"coeff": coeff, "origin": np.array([k0, j0, i0]), "n": np.array([n.value]),
          "in_block": arr[k0 - 1:k0 + 17, j0 - 1:j0 + 17, i0 - 1:i0 + 17].copy(),
          "out_block": out[k0:k0 + 16, j0:j0 + 16, i0:i0 + 16].copy(),
          # first computed corner: its halo lies in the ghost zone of the reference's array
          "in_corner": arr[f - 1:f + 9, f - 1:f + 9, f - 1:f + 9].copy(),
          "out_corner": out[f:f + 8, f:f + 8, f:f + 8].copy()}
    lib.ref3axis_release()
    return fx


D4_DECOMP = [((16, 16, 16, 16), 8), ((20, 20, 20, 20), 8), ((24, 16, 32, 20), 8), ((12, 20, 12, 16), 4),
             ((28, 20, 12, 36), 4)]
D4_RANK_CASES = [((1, 1, 1, 1), (0, 0, 0, 0)), ((2, 1, 1, 1), (1, 0, 0, 0)), ((2, 2, 2, 1), (0, 1, 1, 0)),
                 ((2, 2, 2, 2), (1, 0, 1, 1)), ((3, 2, 1, 2), (2, 1, 0, 1))]
D4_STENCIL_DIMS = [(16, 16, 16, 16), (16, 24, 16, 20)]


def d4_fixtures():
    """The 4-D path of weak/main-4d.cpp from oracle/_ref/libref4d*.so: BrickDecomp<4,4,4,4,4> tables in
    both numberings (page-aligned, -DDECOMP_PAGEUNALIGN) and the generated mpi9pt.py kernel."""
    from oracle import oracle4d as D
    recs = []
    for dims, gz in D4_DECOMP:
        for pa in (True, False):
            ref = D.Ref4D(dims, gz=gz, page_align=pa)
            rec = {"dims": list(dims), "gz": gz, "page_align": pa, "nbricks": int(ref.nbricks),
                   "sep_pos": [int(x) for x in ref.sep_pos], "skin_size": [int(x) for x in ref.skin_size],
                   "adj_fnv": fnv1a64(ref.adj), "grid_fnv": fnv1a64(ref.grid)}
            for name in ("ghost", "skin"):
                rec[name] = {k: [int(x) for x in v] for k, v in ref.regions[name].items()}
            if pa and gz == 8 and dims == D4_DECOMP[0][0]:
                rec["rank_map"] = [{"procdims": list(pd), "coo": list(coo),
                                    "map": {"%x" % k: v for k, v in sorted(ref.rank_map(pd, coo).items())}}
                                   for pd, coo in D4_RANK_CASES]
            ref.close()
            recs.append(rec)
    with open(os.path.join(GOLDEN, "decomp4d.json"), "w") as f:
        json.dump({"generator": "oracle/gen_golden.py d4", "source": "oracle/_ref/libref4d.so, libref4d_unalign.so",
                   "decomp": recs}, f, indent=0, separators=(",", ":"))
    for dims in D4_STENCIL_DIMS:
        ref = D.Ref4D(dims, page_align=False)
        nb = ref.nbricks
        inp = synth.uniform(nb * 256, seed=17)
        out = {"dims": np.array(dims)}
        for s in (0, 1, 2):  # the three sweep regions of one exchange period (skip // TILE)
            blist = D.sweep_list(ref.grid, s)
            res = np.zeros(nb * 256)
            D.ref_brick_sweep(ref.adj, blist, inp, res)
            pick = blist[::max(1, len(blist) // 64)]  # a sample keeps the fixture small
            out["n_%d" % s] = np.array([len(blist)])
            out["bricks_%d" % s] = pick
            out["brick_%d" % s] = res.reshape(nb, 256)[pick].copy()
            untouched = np.setdiff1d(np.arange(nb), blist)
            assert not res.reshape(nb, 256)[untouched].any()  # nothing outside the list is written
        ref.close()
        np.savez_compressed(os.path.join(GOLDEN, "d4pt9_%dx%dx%dx%d.npz" % dims), **out)


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    if "d4" in sys.argv[1:]:  # the 4-D path (weak/main-4d.cpp)
        d4_fixtures()
        for fn in sorted(os.listdir(GOLDEN)):
            if "4d" in fn or "d4" in fn:
                print(fn, os.path.getsize(os.path.join(GOLDEN, fn)))
        return
    if "d3pt27" in sys.argv[1:]:  # the 27-point block (one ~20 s run of the reference's d3pt27())
        np.savez(os.path.join(GOLDEN, "d3pt27_block.npz"), **d3pt27_fixture())
        print("d3pt27_block.npz", os.path.getsize(os.path.join(GOLDEN, "d3pt27_block.npz")))
        return
    if "cond" in sys.argv[1:]:  # only the cond.py fixtures (added after the others were frozen)
        for d in STENCIL_DIMS:
            np.savez(os.path.join(GOLDEN, "cond_%dx%dx%d.npz" % d), **cond_fixture(d))
            print("cond_%dx%dx%d.npz" % d)
        return
    recs = [decomp_record(d) for d in DECOMP_DIMS]
    meta = {"generator": "oracle/gen_golden.py", "source": "oracle/_ref/libref_%s.so" % R.host_isa(),
            "vsvec": R.vsvec(), "decomp": recs}
    with open(os.path.join(GOLDEN, "decomp.json"), "w") as f:
        json.dump(meta, f, indent=0, separators=(",", ":"))
    for d in STENCIL_DIMS:
        fx = stencil_fixture(d)
        np.savez(os.path.join(GOLDEN, "stencil_%dx%dx%d.npz" % d), **fx)
    for fn in sorted(os.listdir(GOLDEN)):
        print(fn, os.path.getsize(os.path.join(GOLDEN, fn)))


if __name__ == "__main__":
    main()
