"""Host-side mirror of ORB_SLAM2::ORBmatcher's distance semantics (reference include/ORBmatcher.h:40-230)
over the C ABI: DescriptorDistance (src/ORBmatcher.cc:2073-2093) and the best/second-best + ratio idiom
(src/ORBmatcher.cc:304-342) as a brute-force k=2 scan on the GPU. No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KNN_DTYPE, MEM_DEVICE, MEM_HOST, check


def _vp(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def hamming_knn(query: np.ndarray, db: np.ndarray, index_base: int = 0, sentinel: int = 256) -> np.ndarray:
    """query [Q,32] u8, db [D,32] u8 host arrays -> KNN_DTYPE[Q] (d1, d2, i1, i2); i = -1 / d = sentinel where
    the reference loop would have left best/second at their initial value."""
    query = np.ascontiguousarray(query, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    if query.ndim != 2 or query.shape[1] != 32 or db.ndim != 2 or db.shape[1] != 32:
        raise TypeError("descriptors must be [N,32] uint8 (256-bit ORB)")
    out = np.zeros(len(query), KNN_DTYPE)
    if len(query) == 0:
        return out
    check(_lib.lib().orb_hamming_knn(_vp(query), MEM_HOST, len(query), _vp(db) if len(db) else None, MEM_HOST, len(db),
                                     index_base, sentinel, _vp(out), MEM_HOST, None))
    return out


def hamming_knn_device(query, db, out=None, index_base: int = 0, sentinel: int = 256, stream=None):
    """torch CUDA tensors: query [Q,32] u8, db [D,32] u8 -> out [Q,3] int64 viewable as orb_knn2 (24 B records).
    Asynchronous on `stream` (torch.cuda.Stream) or the current stream."""
    import torch
    assert query.is_cuda and db.is_cuda and query.dtype == torch.uint8 and db.dtype == torch.uint8
    assert query.is_contiguous() and db.is_contiguous()
    Q, D = query.shape[0], db.shape[0]
    if out is None:
        out = torch.empty((Q, 3), dtype=torch.int64, device=query.device)
    st = (stream or torch.cuda.current_stream(query.device)).cuda_stream
    check(_lib.lib().orb_hamming_knn(query.data_ptr(), MEM_DEVICE, Q, db.data_ptr(), MEM_DEVICE, D, index_base, sentinel,
                                     out.data_ptr(), MEM_DEVICE, C.c_void_p(st)))
    return out


def knn_merge_device(shards, nshards: int, Q: int, out=None, sentinel: int = 256, stream=None):
    """shards: torch int64 CUDA tensor [nshards, Q, 3] of orb_knn2 records in ascending shard order."""
    import torch
    if out is None:
        out = torch.empty((Q, 3), dtype=torch.int64, device=shards.device)
    st = (stream or torch.cuda.current_stream(shards.device)).cuda_stream
    check(_lib.lib().orb_hamming_knn_merge(shards.data_ptr(), nshards, Q, sentinel, out.data_ptr(), C.c_void_p(st)))
    return out


def knn_records(t) -> np.ndarray:
    """torch [Q,3] int64 (device or host) -> KNN_DTYPE[Q] numpy records."""
    return t.detach().cpu().numpy().view(KNN_DTYPE).reshape(-1)


_OWN_COMMS = {}   # (group id, device) -> communicator made through orb_nccl_comm_create


def nccl_comm_ptr(group=None) -> int:
    """The ncclComm_t (as an integer) the C ABI's sharded scan runs its all-gather on: torch.distributed's own communicator of `group`
    (ProcessGroupNCCL._comm_ptr, created lazily by the first collective), or - where this torch does not expose it - one made by
    orb_nccl_comm_create from an id broadcast through `group`."""
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device())
    pg = group if group is not None else dist.distributed_c10d._get_default_group()
    try:
        be = pg._get_backend(dev)
        if hasattr(be, "_comm_ptr"):
            ptr = be._comm_ptr()
            if not ptr:
                dist.all_reduce(torch.zeros(1, device=dev), group=group)   # first collective creates the communicator
                torch.cuda.synchronize(dev)
                ptr = be._comm_ptr()
            if ptr:
                return int(ptr)
    except Exception:   # noqa: BLE001 - fall through to a communicator of our own
        pass
    key = (id(pg), dev.index)
    if key not in _OWN_COMMS:
        L = _lib.lib()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        ident = np.zeros(128, np.uint8)
        if rank == 0:
            check(L.orb_nccl_unique_id(_vp(ident)))
        t = torch.from_numpy(ident).to(dev)
        dist.broadcast(t, src=dist.get_global_rank(pg, 0) if group is not None else 0, group=group)
        ident = t.cpu().numpy()
        comm = C.c_void_p()
        check(L.orb_nccl_comm_create(_vp(ident), world, rank, C.byref(comm)))
        _OWN_COMMS[key] = comm.value
    return _OWN_COMMS[key]


def hamming_knn_sharded(query, db_shard, shard_base: int, group=None, sentinel: int = 256, out=None, comm: int | None = None, stream=None):
    """Database sharded contiguously over the ranks of `group` (ascending rank = ascending rows): orb_hamming_knn_sharded = local k=2 on this
    rank's shard, ONE ncclAllGather of the per-query records, device top-2 merge - all behind the C ABI, asynchronous on the stream.
    Equals one sequential scan over the concatenated database (earliest global index wins ties). query / db_shard: torch uint8 CUDA tensors."""
    import torch
    if not query.is_cuda:
        raise RuntimeError("hamming_knn_sharded needs CUDA tensors (there is no CPU fallback)")
    assert db_shard.is_cuda and query.is_contiguous() and db_shard.is_contiguous()
    Q = query.shape[0]
    if out is None:
        out = torch.empty((Q, 3), dtype=torch.int64, device=query.device)
    if comm is None:
        comm = nccl_comm_ptr(group)
    st = (stream or torch.cuda.current_stream(query.device)).cuda_stream
    check(_lib.lib().orb_hamming_knn_sharded(query.data_ptr(), MEM_DEVICE, Q, db_shard.data_ptr(), MEM_DEVICE, db_shard.shape[0], shard_base, sentinel,
                                             out.data_ptr(), MEM_DEVICE, C.c_void_p(comm), C.c_void_p(st)))
    return out


class KnnPeerGroup:
    """The sharded scan with the exchange inside the kernel (orb_hamming_knn_peer): every rank's reduction kernel stores its two best keys per query
    into the peers' mailboxes over NVLink and merges what the peers stored into its own - no NCCL call on the data path. One process per GPU of one
    node; the mailbox handles (64 bytes per rank) are exchanged once, over `group` (any torch.distributed backend)."""

    def __init__(self, max_q: int, group=None):
        import torch
        import torch.distributed as dist
        self.rank, self.nranks = dist.get_rank(group), dist.get_world_size(group)
        self._h = C.c_void_p()
        # Every rank walks through the same collectives whatever fails locally (a rank that raised early would leave the others hanging in the
        # all-gather): the handle travels with a status byte, and the group only stands if every rank created AND connected.
        err = None
        mine = np.zeros(65, np.uint8)
        try:
            check(_lib.lib().orb_knn_peer_create(self.rank, self.nranks, int(max_q), C.byref(self._h)))
            check(_lib.lib().orb_knn_peer_export(self._h, mine.ctypes.data_as(C.c_void_p)))
            mine[64] = 1
        except Exception as e:   # noqa: BLE001
            err = e
        on_gpu = dist.get_backend(group) == "nccl"
        t = torch.from_numpy(mine)
        if on_gpu:
            t = t.cuda()
        parts = [torch.empty_like(t) for _ in range(self.nranks)]
        dist.all_gather(parts, t, group=group)
        got = np.stack([x.cpu().numpy() for x in parts])
        ok = err is None and bool(got[:, 64].all())
        if ok:
            try:
                handles = np.ascontiguousarray(got[:, :64])
                check(_lib.lib().orb_knn_peer_connect(self._h, handles.ctypes.data_as(C.c_void_p)))
            except Exception as e:   # noqa: BLE001
                err, ok = e, False
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32)
        if on_gpu:
            flag = flag.cuda()
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)   # also: every mailbox is open everywhere before the first store
        if int(flag.item()) == 0:
            self.close()
            raise RuntimeError(f"KnnPeerGroup: not every rank could set up its mailbox ({err if err is not None else 'a peer failed'})")

    def knn(self, query, db_shard, shard_base: int, sentinel: int = 256, out=None, stream=None):
        """Collective. query / db_shard: torch uint8 CUDA tensors; returns out [Q, 3] int64 (orb_knn2 records) as hamming_knn_sharded does."""
        import torch
        assert query.is_cuda and db_shard.is_cuda and query.is_contiguous() and db_shard.is_contiguous()
        Q = query.shape[0]
        if out is None:
            out = torch.empty((Q, 3), dtype=torch.int64, device=query.device)
        st = (stream or torch.cuda.current_stream(query.device)).cuda_stream
        check(_lib.lib().orb_hamming_knn_peer(self._h, query.data_ptr(), MEM_DEVICE, Q, db_shard.data_ptr(), MEM_DEVICE, db_shard.shape[0], shard_base,
                                              sentinel, out.data_ptr(), MEM_DEVICE, C.c_void_p(st)))
        return out

    def check(self, stream=None):
        import torch
        st = (stream or torch.cuda.current_stream()).cuda_stream
        check(_lib.lib().orb_knn_peer_check(self._h, C.c_void_p(st)))

    def close(self):
        if self._h:
            _lib.lib().orb_knn_peer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:   # noqa: BLE001
            pass


class ORBmatcher:
    """include/ORBmatcher.h: constructor (nnratio=0.6, checkOri=true) :54, TH_LOW/TH_HIGH/HISTO_LENGTH :223-225."""
    TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True):
        self.mfNNratio = np.float32(nnratio)
        self.mbCheckOrientation = checkOri

    @staticmethod
    def DescriptorDistance(a: np.ndarray, b: np.ndarray) -> int:
        """src/ORBmatcher.cc:2073-2093 for one pair of 32-byte rows (computed on the GPU)."""
        import torch
        ta = torch.from_numpy(np.ascontiguousarray(a, np.uint8).reshape(1, 32)).cuda()
        tb = torch.from_numpy(np.ascontiguousarray(b, np.uint8).reshape(1, 32)).cuda()
        return int(ORBmatcher.DescriptorDistanceBatch(ta, tb)[0])

    @staticmethod
    def DescriptorDistanceBatch(a, b):
        """n row pairs on the device: torch uint8 [n,32] x2 -> int32 [n]."""
        import torch
        assert a.is_cuda and b.is_cuda and a.shape == b.shape and a.is_contiguous() and b.is_contiguous()
        out = torch.empty((a.shape[0],), dtype=torch.int32, device=a.device)
        st = torch.cuda.current_stream(a.device).cuda_stream
        check(_lib.lib().orb_descriptor_distance_batch(a.data_ptr(), b.data_ptr(), a.shape[0], out.data_ptr(), C.c_void_p(st)))
        return out

    def accept(self, knn: np.ndarray, th: int | None = None) -> np.ndarray:
        """The accept test of src/ORBmatcher.cc:339-342: best <= TH_LOW and (float)best < mfNNratio*(float)second."""
        th = self.TH_LOW if th is None else th
        return (knn["d1"] <= th) & (knn["d1"].astype(np.float32) < self.mfNNratio * knn["d2"].astype(np.float32))

    def match(self, query: np.ndarray, db: np.ndarray, th: int | None = None):
        """Brute-force k=2 + ratio test. Returns (knn records, accept mask)."""
        knn = hamming_knn(query, db)
        return knn, self.accept(knn, th)

    # ---- guided matching ----
    def SearchByProjection(self, kps_un: np.ndarray, desc: np.ndarray, n_kps: np.ndarray, queries: np.ndarray, qdesc: np.ndarray,
                           n_queries: np.ndarray, bounds, th_dist: int | None = None, use_ratio: bool = False, uright=None,
                           claimed=None):
        """The candidate search of ORBmatcher::SearchByProjection for `frames` independent (query set, target frame) pairs.

        use_ratio=False : Frame-to-Frame tracking form (src/ORBmatcher.cc:1702-1866): best only, rotation histogram iff
                          mbCheckOrientation.
        use_ratio=True  : local-map form (src/ORBmatcher.cc:69-184): best / second best with the level-conditional
                          mfNNratio test, no rotation histogram.
        kps_un [F,cap] KP_DTYPE (undistorted keypoints of the target frames), desc [F,cap,32], n_kps [F];
        queries [F,qcap] QUERY_DTYPE (projection, radius, level window, flags), qdesc [F,qcap,32], n_queries [F];
        bounds = (mnMinX, mnMinY, mnMaxX, mnMaxY); uright [F,cap] float32 (mvuRight) or None; claimed [F,cap] uint8 or None.
        Returns (assigned [F,cap] int32: query index per keypoint or -1, query_match [F,qcap] int32, nmatches [F])."""
        from ._lib import QUERY_DTYPE, KP_DTYPE, OrbSearchParams
        F, cap = kps_un.shape
        qcap = queries.shape[1]
        assert kps_un.dtype == KP_DTYPE and queries.dtype == QUERY_DTYPE and queries.shape[0] == F
        assert desc.shape == (F, cap, 32) and qdesc.shape == (F, qcap, 32) and desc.dtype == np.uint8 and qdesc.dtype == np.uint8
        arrs = [np.ascontiguousarray(a) for a in (kps_un, desc, queries, qdesc)]
        nk = np.ascontiguousarray(n_kps, np.int32)
        nq = np.ascontiguousarray(n_queries, np.int32)
        ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
        cl = None if claimed is None else np.ascontiguousarray(claimed, np.uint8)
        prm = OrbSearchParams(float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]),
                              int(self.TH_HIGH if th_dist is None else th_dist), float(self.mfNNratio) if use_ratio else 0.0,
                              int(self.mbCheckOrientation and not use_ratio), 0)
        assigned = np.empty((F, cap), np.int32)
        qmatch = np.empty((F, qcap), np.int32)
        nm = np.empty(F, np.int32)
        check(_lib.lib().orb_search_by_projection(MEM_HOST, _vp(arrs[0]), _vp(arrs[1]), None if ur is None else _vp(ur),
                                                  None if cl is None else _vp(cl), _vp(nk), cap, _vp(arrs[2]), _vp(arrs[3]), _vp(nq),
                                                  qcap, F, C.byref(prm), _vp(assigned), _vp(qmatch), _vp(nm), None))
        return assigned, qmatch, nm

    @staticmethod
    def ComputeDistinctiveDescriptors(desc: np.ndarray, offsets: np.ndarray):
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517) for many map points at once: desc rows
        [offsets[p], offsets[p+1]) are the observations of point p. Returns (best_idx [P] relative row or -1, best_median [P])."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        offsets = np.ascontiguousarray(offsets, np.int64)
        P = len(offsets) - 1
        bi, bm = np.empty(max(P, 1), np.int32), np.empty(max(P, 1), np.int32)
        check(_lib.lib().orb_distinctive_descriptors(MEM_HOST, _vp(desc) if len(desc) else None, _vp(offsets), P, _vp(bi), _vp(bm), None))
        return bi[:P], bm[:P]

    def SearchByBoW(self, kf_desc, kf_kps_un, kf_valid, kf_bow: dict, f_desc, f_kps, f_bow: dict, th_dist: int | None = None, f_blocked=None):
        """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:248-410) for F independent pairs.
        kf_desc [F,qcap,32], kf_kps_un [F,qcap], kf_valid [F,qcap] (feature has a good map point), f_desc [F,cap,32], f_kps [F,cap];
        kf_bow / f_bow = the dicts returned by ORBVocabulary.transform for the same arrays. f_blocked [F,cap] (optional) marks target
        features that may not be matched: with it and th_dist = TH_LOW - 1 this is the KeyFrame-KeyFrame form (:722-880).
        Returns (assigned [F,cap], kf_match [F,qcap], nmatches [F])."""
        F, qcap = kf_kps_un.shape
        cap = f_kps.shape[1]
        c = np.ascontiguousarray
        args = [c(kf_desc, np.uint8), c(kf_kps_un), c(kf_valid, np.uint8), c(kf_bow["fv_node"]), c(kf_bow["fv_start"]), c(kf_bow["fv_feature"]),
                c(kf_bow["fv_count"]), c(f_desc, np.uint8), c(f_kps), None if f_blocked is None else c(f_blocked, np.uint8), c(f_bow["fv_node"]),
                c(f_bow["fv_start"]), c(f_bow["fv_feature"]), c(f_bow["fv_count"])]
        assert args[4].shape == (F, qcap + 1) and args[11].shape == (F, cap + 1)
        assigned = np.empty((F, cap), np.int32)
        kf_match = np.empty((F, qcap), np.int32)
        nm = np.empty(F, np.int32)
        p = [None if a is None else _vp(a) for a in args]
        check(_lib.lib().orb_search_by_bow(MEM_HOST, *p[:7], qcap, *p[7:], cap, F, int(self.TH_LOW if th_dist is None else th_dist),
                                           float(self.mfNNratio), int(self.mbCheckOrientation), _vp(assigned), _vp(kf_match), _vp(nm), None))
        return assigned, kf_match, nm

    def FuseSearch(self, kps_un, desc, n_kps, queries, qdesc, n_queries, bounds, uright=None, inv_level_sigma2=None, th_dist: int | None = None):
        """The candidate search of ORBmatcher::Fuse (src/ORBmatcher.cc:1100-1280, 1282-1430) and of SearchBySim3's two projection passes
        (:1433-1700) for F keyframes: every projected map point on its own (no claims). queries [F,qcap] QUERY_DTYPE: u, v, radius,
        ur, level window (min_level, max_level) = (nPredictedLevel - 1, nPredictedLevel). inv_level_sigma2 (mvInvLevelSigma2) switches
        the chi-square reprojection test of Fuse(pKF, vpMapPoints, th) on. Returns (best_idx [F,qcap] or -1, best_dist [F,qcap])."""
        from ._lib import OrbSearchParams
        F, cap = kps_un.shape
        qcap = queries.shape[1]
        c = np.ascontiguousarray
        a = [c(kps_un), c(desc, np.uint8), c(n_kps, np.int32), c(queries), c(qdesc, np.uint8), c(n_queries, np.int32)]
        ur = None if uright is None else c(uright, np.float32)
        sig = None
        if inv_level_sigma2 is not None:
            sig = np.zeros(16, np.float32)
            sig[:len(inv_level_sigma2)] = inv_level_sigma2
        prm = OrbSearchParams(float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), int(self.TH_LOW if th_dist is None else th_dist),
                              0.0, 0, 0)
        bi, bd = np.empty((F, qcap), np.int32), np.empty((F, qcap), np.int32)
        check(_lib.lib().orb_fuse_search(MEM_HOST, _vp(a[0]), _vp(a[1]), None if ur is None else _vp(ur), _vp(a[2]), cap, _vp(a[3]), _vp(a[4]),
                                         _vp(a[5]), qcap, F, C.byref(prm), None if sig is None else _vp(sig), _vp(bi), _vp(bd), None))
        return bi, bd

    def SearchForTriangulation(self, kf1_desc, kf1_kps_un, kf1_has_mp, kf1_uright, kf1_bow: dict, kf2_desc, kf2_kps_un, kf2_has_mp, kf2_uright,
                               kf2_bow: dict, epi, only_stereo: bool = False, th_dist: int | None = None):
        """ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:879-1090) for P independent
        keyframe pairs. kf?_desc [P,n,32], kf?_kps_un [P,n] (mvKeysUn), kf?_has_mp [P,n] (GetMapPoint(idx) != NULL), kf?_uright [P,n]
        (mvuRight), kf?_bow = ORBVocabulary.transform output; epi [P] EPI_DTYPE (F12, epipole, pKF2's sigma2 / scale tables).
        Returns (vMatches12 [P,qcap], nmatches [P]); vMatchedPairs of pair p = [(i, m) for i, m in enumerate(vMatches12[p]) if m >= 0]."""
        from ._lib import EPI_DTYPE
        P, qcap = kf1_kps_un.shape
        cap = kf2_kps_un.shape[1]
        c = np.ascontiguousarray
        ur1, ur2 = c(kf1_uright, np.float32), c(kf2_uright, np.float32)
        free1 = (c(kf1_has_mp, np.uint8) == 0)
        blocked2 = (c(kf2_has_mp, np.uint8) != 0)
        if only_stereo:   # :931-934, :963-966
            free1 &= ur1 >= 0
            blocked2 |= ~(ur2 >= 0)
        epi = c(epi)
        assert epi.dtype == EPI_DTYPE and epi.shape == (P,)
        args = [c(kf1_desc, np.uint8), c(kf1_kps_un), c(free1, np.uint8), ur1, c(kf1_bow["fv_node"]), c(kf1_bow["fv_start"]), c(kf1_bow["fv_feature"]),
                c(kf1_bow["fv_count"]), c(kf2_desc, np.uint8), c(kf2_kps_un), c(blocked2, np.uint8), ur2, c(kf2_bow["fv_node"]),
                c(kf2_bow["fv_start"]), c(kf2_bow["fv_feature"]), c(kf2_bow["fv_count"])]
        assert args[5].shape == (P, qcap + 1) and args[13].shape == (P, cap + 1)
        m12, nm = np.empty((P, qcap), np.int32), np.empty(P, np.int32)
        p = [_vp(a) for a in args]
        check(_lib.lib().orb_search_for_triangulation(MEM_HOST, *p[:8], qcap, *p[8:], cap, P, _vp(epi), int(self.TH_LOW if th_dist is None else th_dist),
                                                      int(self.mbCheckOrientation), _vp(m12), _vp(nm), None))
        return m12, nm

    def SearchForInitialization(self, kps2_un, desc2, n2, queries, qdesc, n_queries, bounds, th_dist: int | None = None):
        """ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:570-712) for F frame
        pairs. queries [F,qcap] QUERY_DTYPE: u, v = vbPrevMatched, radius = windowSize, levels (0, 0), angle, SKIP for level > 0.
        Returns (vnMatches12 [F,qcap], matches21 [F,cap], nmatches [F])."""
        from ._lib import OrbSearchParams
        F, cap = kps2_un.shape
        qcap = queries.shape[1]
        c = np.ascontiguousarray
        a = [c(kps2_un), c(desc2, np.uint8), c(n2, np.int32), c(queries), c(qdesc, np.uint8), c(n_queries, np.int32)]
        prm = OrbSearchParams(float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), int(self.TH_LOW if th_dist is None else th_dist),
                              float(self.mfNNratio), int(self.mbCheckOrientation), 0)
        m12, m21, nm = np.empty((F, qcap), np.int32), np.empty((F, cap), np.int32), np.empty(F, np.int32)
        check(_lib.lib().orb_search_for_initialization(MEM_HOST, _vp(a[0]), _vp(a[1]), _vp(a[2]), cap, _vp(a[3]), _vp(a[4]), _vp(a[5]), qcap, F,
                                                       C.byref(prm), _vp(m12), _vp(m21), _vp(nm), None))
        return m12, m21, nm
