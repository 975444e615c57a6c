"""world_size-2 run of the N>1 host logic on CPU (gloo): two processes execute the product's
exchange plans -- the per-region weak plan of BrickDecomp (brick-mpi.h:466-495), its SHIFT form
(brick-mpi.h:587-681) and the per-peer
packed strong plan (strong/main.cu:188-443) -- with torch.distributed messages on host buffers,
and the result must equal the oracle's multi-rank emulation bit for bit."""
import os
import socket
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_exchange_plans_over_gloo(built, oracle, tmp_path):
    world, port = 2, _free_port()
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "gloo_worker.py"), str(r), str(world), str(port),
                               str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(world)]
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out.decode()[-3000:]

    # weak: 2 ranks on the periodic process grid MPI_Dims_create gives (2,1,1)
    dims = (32, 16, 16)
    odec = oracle.Decomp(dims, 8)
    ins = [np.load(tmp_path / ("weak_in_%d.npy" % r)) for r in range(world)]
    exp = oracle.exchange([x.copy() for x in ins], odec, built.cart_dims(world))
    for r in range(world):
        got = np.load(tmp_path / ("weak_out_%d.npy" % r))
        assert np.array_equal(got.view(np.uint64), exp[r].view(np.uint64)), r
        assert not np.array_equal(got, ins[r])
        # the SHIFT exchange (three stages, six messages) leaves the same storage
        shifted = np.load(tmp_path / ("shift_out_%d.npy" % r))
        assert np.array_equal(shifted.view(np.uint64), exp[r].view(np.uint64)), r

    # strong: 8 subdomains of 32^3, Morton ids 0-3 on rank 0 and 4-7 on rank 1
    sdec = oracle.Decomp((32, 32, 32), 8)
    sin = {i: np.load(tmp_path / ("strong_in_%d.npy" % i)) for i in range(8)}
    sexp = oracle.strong_exchange({k: v.copy() for k, v in sin.items()}, sdec, 2)
    for i in range(8):
        got = np.load(tmp_path / ("strong_out_%d.npy" % i))
        assert np.array_equal(got.view(np.uint64), sexp[i].view(np.uint64)), i

    # 4-D (weak/main-4d.cpp): 2 ranks on the (2,1,1,1) grid, 312 regions each
    from oracle import oracle4d as D
    ddec = D.DecompND((16, 16, 16, 16), page_align=False)
    dins = [np.load(tmp_path / ("weak4d_in_%d.npy" % r)) for r in range(world)]
    dexp = [x.copy() for x in dins]
    D.exchange(dexp, ddec, built.cart_dims(world, 4))
    for r in range(world):
        got = np.load(tmp_path / ("weak4d_out_%d.npy" % r))
        assert np.array_equal(got.view(np.uint64), dexp[r].view(np.uint64)), r
        assert not np.array_equal(got, dins[r])
