"""Worker for tests/test_gloo_world2.py: one rank of a world_size-2 gloo group that runs the
product's host-side exchange plans (weak: bricks_decomp_exchange_plan; strong: StrongPlan) with
torch.distributed point-to-point messages on CPU buffers in place of NCCL on device buffers."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _exchange_bytes(dist, torch, rank, buf, recvs, sends):
    """recvs/sends: lists of (peer, offset, nbytes, tag) on the uint8 tensor `buf`."""
    reqs, keep = [], []
    for peer, off, n, tag in recvs:
        if peer != rank and n:
            t = torch.empty(n, dtype=torch.uint8)
            keep.append((t, off, n))
            reqs.append(dist.irecv(t, src=peer, tag=tag))
    for peer, off, n, tag in sends:
        if peer != rank and n:
            reqs.append(dist.isend(buf[off:off + n].clone(), dst=peer, tag=tag))
    # messages to myself: plain copies, paired by tag
    mine = {tag: (off, n) for peer, off, n, tag in sends if peer == rank}
    local = [(off, mine[tag][0], n) for peer, off, n, tag in recvs if peer == rank and n]
    src = buf.clone()
    for doff, soff, n in local:
        buf[doff:doff + n] = src[soff:soff + n]
    for r in reqs:
        r.wait()
    for t, off, n in keep:
        buf[off:off + n] = t


def weak(rank, world, outdir):
    import torch
    import torch.distributed as dist
    from bricklib_b200 import capi
    dims = (32, 16, 16)
    procdims = capi.cart_dims(world)
    dec = capi.Decomp(dims, 8)
    dec.rank_map(procdims, capi.cart_coords(rank, procdims))
    ro, so, nb, rp, sp = dec.exchange_plan(512)
    field = capi.uniform(dec.nbricks * 512, seed=500 + rank)
    np.save(os.path.join(outdir, "weak_in_%d.npy" % rank), field)
    buf = torch.from_numpy(field.view(np.uint8).copy())
    n = len(ro)
    _exchange_bytes(dist, torch, rank, buf,
                    [(int(rp[i]), int(ro[i]), int(nb[i]), i) for i in range(n)],
                    [(int(sp[i]), int(so[i]), int(nb[i]), i) for i in range(n)])
    np.save(os.path.join(outdir, "weak_out_%d.npy" % rank), buf.numpy().view(np.float64))


def shift(rank, world, outdir):
    """the SHIFT exchange (bricks_decomp_shift_plan): three stages in order, piece i of a stage is
    piece i on both ends (both ranks hold the same tables), so the piece index is the tag"""
    import torch
    import torch.distributed as dist
    from bricklib_b200 import capi
    dims = (32, 16, 16)
    procdims = capi.cart_dims(world)
    dec = capi.Decomp(dims, 8)
    dec.rank_map(procdims, capi.cart_coords(rank, procdims))
    field = capi.uniform(dec.nbricks * 512, seed=500 + rank)  # same input as weak()
    buf = torch.from_numpy(field.view(np.uint8).copy())
    for stage, (ro, so, nb, rp, sp) in enumerate(dec.shift_plan(512)):
        n = len(ro)
        _exchange_bytes(dist, torch, rank, buf,
                        [(int(rp[i]), int(ro[i]), int(nb[i]), 1000 * stage + i) for i in range(n)],
                        [(int(sp[i]), int(so[i]), int(nb[i]), 1000 * stage + i) for i in range(n)])
        dist.barrier()
    np.save(os.path.join(outdir, "shift_out_%d.npy" % rank), buf.numpy().view(np.float64))


def strong(rank, world, outdir):
    import torch
    import torch.distributed as dist
    from bricklib_b200 import capi
    dom, sdom = 64, 32
    dec = capi.Decomp((sdom,) * 3, 8)
    plan = capi.StrongPlan(dom, sdom, world, rank, dec)
    per = dec.nbricks * 512 * 8
    fields = [capi.uniform(dec.nbricks * 512, seed=900 + plan.first_id + s) for s in range(plan.nsub)]
    for s, f in enumerate(fields):
        np.save(os.path.join(outdir, "strong_in_%d.npy" % (plan.first_id + s)), f)
    store = torch.from_numpy(np.concatenate(fields).view(np.uint8).copy())
    # pack -> one message per peer -> unpack -> local links (strong/main.cu:397-443)
    reqs, inbox = [], []
    for peer in plan.recv:
        t = torch.empty(peer["bytes"], dtype=torch.uint8)
        inbox.append((peer, t))
        reqs.append(dist.irecv(t, src=peer["rank"], tag=0))
    for peer in plan.send:
        msg = torch.empty(peer["bytes"], dtype=torch.uint8)
        for sub, _, off, n, boff, _ in peer["pieces"]:
            msg[boff:boff + n] = store[sub * per + off:sub * per + off + n]
        reqs.append(dist.isend(msg, dst=peer["rank"], tag=0))
    for r in reqs:
        r.wait()
    src = store.clone()
    for peer, t in inbox:
        for sub, _, off, n, boff, _ in peer["pieces"]:
            store[sub * per + off:sub * per + off + n] = t[boff:boff + n]
    for ss, ds, so, do, n in plan.links:
        store[ds * per + do:ds * per + do + n] = src[ss * per + so:ss * per + so + n]
    out = store.numpy().view(np.float64)
    for s in range(plan.nsub):
        np.save(os.path.join(outdir, "strong_out_%d.npy" % (plan.first_id + s)), out[s * per // 8:(s + 1) * per // 8])


def weak4d(rank, world, outdir):
    """the 4-D decomposition (weak/main-4d.cpp): 312 regions per rank on the process grid MPI_Dims_create
    gives for 4 axes; region i is message i on both ends"""
    import torch
    import torch.distributed as dist
    from bricklib_b200 import capi
    dims = (16, 16, 16, 16)
    procdims = capi.cart_dims(world, 4)
    dec = capi.Decomp4D(dims, 8, False)
    dec.rank_map(procdims, capi.cart_coords(rank, procdims))
    ro, so, nb, rp, sp = dec.exchange_plan(256)
    field = capi.uniform(dec.nbricks * 256, seed=700 + rank)
    np.save(os.path.join(outdir, "weak4d_in_%d.npy" % rank), field)
    buf = torch.from_numpy(field.view(np.uint8).copy())
    n = len(ro)
    _exchange_bytes(dist, torch, rank, buf,
                    [(int(rp[i]), int(ro[i]), int(nb[i]), i) for i in range(n)],
                    [(int(sp[i]), int(so[i]), int(nb[i]), i) for i in range(n)])
    np.save(os.path.join(outdir, "weak4d_out_%d.npy" % rank), buf.numpy().view(np.float64))


def main():
    rank, world, port, outdir = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    weak(rank, world, outdir)
    dist.barrier()
    shift(rank, world, outdir)
    dist.barrier()
    strong(rank, world, outdir)
    dist.barrier()
    weak4d(rank, world, outdir)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
