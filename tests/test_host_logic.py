"""Host-side logic that needs no GPU: level geometry / algorithmic bytes, synthetic generators, shard arithmetic,
the shard-merge rule, and the world_size-2 gloo exchange of the sharded kNN (the oracle stands in for the local scan)."""
import os
import socket
import sys

import numpy as np
import pytest

import knn_merge_ref
from orb_slam2_detailed_comments_b200 import geometry, sharding, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_level_sizes_and_bytes():
    # TUM interior sizes and the algorithmic bytes of every config
    assert [s[:2] for s in geometry.level_sizes(640, 480)] == [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231),
                                                               (257, 193), (214, 161), (179, 134)]
    assert geometry.level_sizes(1241, 376)[0][2:] == (1279, 414)
    assert geometry.algorithmic_bytes(640, 480, 1000) == 6_119_936
    assert geometry.algorithmic_bytes(1241, 376, 2000) == 9_224_499
    assert geometry.algorithmic_bytes(752, 480, 1200) == 7_112_769
    assert geometry.algorithmic_bytes(3840, 2160, 8000) == 141_758_480
    sb = geometry.stage_bytes(640, 480, 1000)
    assert sb["fast_cells"] == 1_158_012 and sb["blur"] == 2 * 1_158_012


def test_scale_factors_match_oracle(oracle):
    assert np.array_equal(geometry.scale_factors(8, 1.2), oracle.OracleExtractor(1000).table(0))


def test_synth_is_deterministic():
    a, b = synth.synth_frame(320, 240, 5), synth.synth_frame(320, 240, 5)
    assert a.dtype == np.uint8 and a.shape == (240, 320) and np.array_equal(a, b)
    assert not np.array_equal(a, synth.synth_frame(320, 240, 6))
    batch = synth.synth_batch(160, 120, 5, seed0=1, distinct=2)
    assert batch.shape == (5, 120, 160) and not np.array_equal(batch[0], batch[2])
    q, db = synth.synth_descriptors(100, 1000, seed=7)
    assert q.shape == (100, 32) and db.shape == (1000, 32)
    d = np.unpackbits(q[:10, None, :] ^ db[None], axis=2).sum(2).min(1)
    assert (d <= 40).all()          # planted queries


def test_shard_ranges():
    for total in (0, 1, 7, 8, 1000, 1_000_000):
        for world in (1, 2, 3, 8):
            rs = [sharding.shard_range(total, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [e - b for b, e in rs]
            assert max(sizes) - min(sizes) <= 1
    assert sharding.db_shard(1_000_000, 3, 8) == (375_000, 500_000)
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def test_merge_rule_equals_sequential_scan(oracle):
    from orb_slam2_detailed_comments_b200 import matcher
    q, db = synth.synth_descriptors(120, 5000, seed=13)
    db[4000:4040] = db[10:50]              # ties across shards: earliest global index must win
    want = oracle.hamming_knn(q, db)
    for bounds in ([0, 5000], [0, 2500, 5000], [0, 1, 1, 4999, 5000], [0, 700, 1400, 2100, 2800, 3500, 4200, 4900, 5000]):
        parts = [oracle.hamming_knn(q, db[a:b], index_base=a) for a, b in zip(bounds[:-1], bounds[1:])]
        got = knn_merge_ref.merge_records_host(parts)
        for f in ("d1", "d2", "i1", "i2"):
            assert np.array_equal(got[f], want[f]), (f, bounds)
    # sentinel handling: nothing below the sentinel anywhere
    ones, zeros = np.full((3, 32), 255, np.uint8), np.zeros((6, 32), np.uint8)
    parts = [oracle.hamming_knn(ones, zeros[:3]), oracle.hamming_knn(ones, zeros[3:], index_base=3)]
    got = knn_merge_ref.merge_records_host(parts)
    assert (got["d1"] == 256).all() and (got["i1"] == -1).all() and (got["i2"] == -1).all()


def test_ratio_accept_expression():
    from orb_slam2_detailed_comments_b200 import ORBmatcher, KNN_DTYPE
    m = ORBmatcher(0.6)
    k = np.zeros(4, KNN_DTYPE)
    k["d1"] = [30, 30, 51, 0]
    k["d2"] = [51, 49, 200, 0]
    # src/ORBmatcher.cc:339-342: best <= TH_LOW and best < 0.6f * second
    assert m.accept(k).tolist() == [True, False, False, False]
    assert (ORBmatcher.TH_LOW, ORBmatcher.TH_HIGH, ORBmatcher.HISTO_LENGTH) == (50, 100, 30)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, q, db, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle_py
    from orb_slam2_detailed_comments_b200 import matcher, sharding as sh
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        b, e = sh.db_shard(len(db), rank, world)
        local = oracle_py.hamming_knn(q, db[b:e], index_base=b)      # checker standing in for the device scan
        t = torch.from_numpy(local.view(np.int64).reshape(len(q), 3).copy())
        gathered = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)                                  # the path's one exchange step
        parts = [g.numpy().reshape(-1).view(matcher.KNN_DTYPE) for g in gathered]
        merged = knn_merge_ref.merge_records_host(parts)
        fb, fe = sh.frame_shard(11, rank, world)
        cnt = torch.tensor([fe - fb])
        dist.all_reduce(cnt)
        if rank == 0:
            ret["merged"] = merged
            ret["frames"] = int(cnt.item())
    finally:
        dist.destroy_process_group()


def test_sharded_knn_world2_gloo(oracle):
    import torch.multiprocessing as mp
    q, db = synth.synth_descriptors(64, 3001, seed=17)
    db[2000:2010] = db[5:15]
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q, db, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    want = oracle.hamming_knn(q, db)
    got = ret["merged"]
    for f in ("d1", "d2", "i1", "i2"):
        assert np.array_equal(got[f], want[f]), f
    assert ret["frames"] == 11
