"""orb_hamming_knn_sharded on 2 GPUs over NCCL: the sharded database scan of config C4 through the C ABI — shard scan,
ncclAllGather of the Q x 24-byte records, device merge — against the oracle's ONE scan over the whole database, cross-shard ties included
(the earliest GLOBAL index must win, src/ORBmatcher.cc:304-335). Needs 2 GPUs (skipped otherwise; `gpurun --gpus 2`)."""
import os
import socket
import sys

import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q, db, bounds, id_path, ret):
    sys.path.insert(0, ROOT)
    import ctypes as C
    import torch
    import torch.distributed as dist
    from orb_slam2_detailed_comments_b200 import _lib, matcher
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        dev = torch.device("cuda", rank)
        b, e = bounds[rank], bounds[rank + 1]
        tq = torch.from_numpy(q).to(dev)
        tdb = torch.from_numpy(db[b:e].copy()).to(dev)
        # (1) torch.distributed's communicator handed to the C ABI
        got1 = matcher.knn_records(matcher.hamming_knn_sharded(tq, tdb, b))
        # (2) a communicator made by the C ABI's own helpers; rank 0's id travels through a file
        L = _lib.lib()
        ident = np.zeros(128, np.uint8)
        if rank == 0:
            _lib.check(L.orb_nccl_unique_id(ident.ctypes.data_as(C.c_void_p)))
            with open(id_path + ".tmp", "wb") as f:
                f.write(ident.tobytes())
            os.replace(id_path + ".tmp", id_path)
        dist.barrier()
        ident = np.frombuffer(open(id_path, "rb").read(), np.uint8).copy()
        comm = C.c_void_p()
        _lib.check(L.orb_nccl_comm_create(ident.ctypes.data_as(C.c_void_p), world, rank, C.byref(comm)))
        out = torch.zeros((len(q), 3), dtype=torch.int64, device=dev)
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        for _ in range(3):   # back to back on a side stream: the shared scratch is reused
            matcher.hamming_knn_sharded(tq, tdb, b, out=out, comm=comm.value, stream=s, sentinel=2**31 - 1)
        s.synchronize()
        got2 = matcher.knn_records(out)
        # (3) host buffers in and out (what a C++ caller with cv::Mat data passes)
        hq, hdb = np.ascontiguousarray(q), np.ascontiguousarray(db[b:e])
        hout = np.zeros(len(q), matcher.KNN_DTYPE)
        _lib.check(L.orb_hamming_knn_sharded(hq.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, len(q), hdb.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, e - b, b, 256,
                                             hout.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, comm, None))
        L.orb_nccl_comm_destroy(comm)
        ret[rank] = (got1, got2, hout)
    finally:
        dist.destroy_process_group()


def test_sharded_knn_two_ranks_nccl(oracle, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    q, db = synth.synth_descriptors(700, 40_001, seed=29)
    db[30_000:30_050] = db[100:150]          # duplicates of shard-0 rows inside shard 1: ties across shards
    q[:50] = db[100:150]                     # ... that are exact matches of queries: d1 = d2 = 0, i1 in shard 0, i2 in shard 1
    bounds = [0, 17_000, 40_001]             # uneven contiguous shards
    want = oracle.hamming_knn(q, db, nthreads=8)
    want_big = oracle.hamming_knn(q, db, sentinel=2**31 - 1, nthreads=8)
    assert (want["i1"][:50] < 17_000).all() and (want["i2"][:50] >= 17_000).all()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), q, db, bounds, str(tmp_path / "nccl_id"), ret), nprocs=2, join=True)
    for rank in range(2):                    # every rank holds the merged result
        got1, got2, hout = ret[rank]
        for f in ("d1", "d2", "i1", "i2"):
            assert np.array_equal(got1[f], want[f]), (rank, "torch comm", f)
            assert np.array_equal(got2[f], want_big[f]), (rank, "own comm", f)
            assert np.array_equal(hout[f], want[f]), (rank, "host buffers", f)


def _peer_worker(rank, world, port, q, db, bounds, ret):
    sys.path.insert(0, ROOT)
    import ctypes as C
    import torch
    import torch.distributed as dist
    from orb_slam2_detailed_comments_b200 import _lib, matcher
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)   # the data path needs no NCCL
    try:
        dev = torch.device("cuda", rank)
        b, e = bounds[rank], bounds[rank + 1]
        tq = torch.from_numpy(q).to(dev)
        tdb = torch.from_numpy(db[b:e].copy()).to(dev)
        grp = matcher.KnnPeerGroup(max_q=1024)
        got1 = matcher.knn_records(grp.knn(tq, tdb, b))
        out = torch.zeros((len(q), 3), dtype=torch.int64, device=dev)
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        for _ in range(5):   # back to back: both parities of the mailbox, epochs, the ticket reset
            grp.knn(tq, tdb, b, out=out, stream=s, sentinel=2**31 - 1)
        grp.check(s)
        got2 = matcher.knn_records(out)
        # fewer queries than max_q, host buffers (synchronous, error check inside the call)
        L = _lib.lib()
        hq, hdb = np.ascontiguousarray(q[:333]), np.ascontiguousarray(db[b:e])
        hout = np.zeros(333, matcher.KNN_DTYPE)
        _lib.check(L.orb_hamming_knn_peer(grp._h, hq.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, 333, hdb.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, e - b, b, 256,
                                          hout.ctypes.data_as(C.c_void_p), _lib.MEM_HOST, None))
        # capacity is refused before anything is launched
        big = torch.zeros((1025, 32), dtype=torch.uint8, device=dev)
        try:
            grp.knn(big, tdb, b)
            refused = False
        except _lib.OrbError:
            refused = True
        dist.barrier()
        grp.close()
        ret[rank] = (got1, got2, hout, refused)
    finally:
        dist.destroy_process_group()


def test_sharded_knn_two_ranks_peer_stores(oracle):
    """orb_hamming_knn_peer: the same sharded scan with the all-gather done by stores into the peers' mailboxes from inside the reduction kernel
    (CUDA IPC, NVLink) and the merge in the same kernel - against the oracle's ONE scan, cross-shard ties included."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    q, db = synth.synth_descriptors(700, 40_001, seed=31)
    db[30_000:30_050] = db[100:150]
    q[:50] = db[100:150]
    bounds = [0, 17_000, 40_001]
    want = oracle.hamming_knn(q, db, nthreads=8)
    want_big = oracle.hamming_knn(q, db, sentinel=2**31 - 1, nthreads=8)
    ret = mp.Manager().dict()
    mp.spawn(_peer_worker, args=(2, _free_port(), q, db, bounds, ret), nprocs=2, join=True)
    for rank in range(2):
        got1, got2, hout, refused = ret[rank]
        assert refused
        for f in ("d1", "d2", "i1", "i2"):
            assert np.array_equal(got1[f], want[f]), (rank, "first call", f)
            assert np.array_equal(got2[f], want_big[f]), (rank, "sixth call, side stream", f)
            assert np.array_equal(hout[f], want[f][:333]), (rank, "host buffers", f)
