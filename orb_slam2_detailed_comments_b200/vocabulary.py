"""ORBVocabulary (include/ORBVocabulary.h:32-33 = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) on the GPU:
loadFromTextFile (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1368-1440) on the host, transform() =
Frame::ComputeBoW (src/Frame.cc:750-765) through the C ABI (orb_vocabulary_create / orb_bow_transform). No CPU path."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import MEM_HOST, check


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def read_text_vocabulary(path: str):
    """Parses the ORBvoc.txt format: header 'k L scoring weighting', then one line per node (ids 1.. in line order):
    'parent is_leaf d0 .. d31 weight'. Returns (k, L, scoring, weighting, parent, is_leaf, desc, weight) with entry 0 = root."""
    with open(path) as f:
        k, L, scoring, weighting = (int(x) for x in f.readline().split())
        rows = [ln.split() for ln in f if ln.strip()]
    n = len(rows) + 1
    parent = np.zeros(n, np.int32)
    is_leaf = np.zeros(n, np.uint8)
    desc = np.zeros((n, 32), np.uint8)
    weight = np.zeros(n, np.float64)
    for i, r in enumerate(rows, start=1):
        parent[i], is_leaf[i] = int(r[0]), int(r[1]) > 0
        desc[i] = [int(x) for x in r[2:34]]
        weight[i] = float(r[34])
    return k, L, scoring, weighting, parent, is_leaf, desc, weight


def write_text_vocabulary(path: str, k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0, trailing_newline=True):
    """saveToTextFile counterpart (TemplatedVocabulary.h:1560-1590): repr() keeps every bit of the double weights.
    trailing_newline=False writes the last node line without '\n': the reference's loadFromTextFile loops on `!f.eof()`
    (TemplatedVocabulary.h:1411) and turns the empty line after a final newline into one more child of the root whose
    descriptor is uninitialised memory; read_text_vocabulary() here skips blank lines (canonicalisation 7)."""
    with open(path, "w") as f:
        f.write(f"{k} {L} {scoring} {weighting}\n")
        lines = (f"{int(parent[i])} {int(is_leaf[i])} " + " ".join(str(int(b)) for b in desc[i]) + f" {float(weight[i])!r}"
                 for i in range(1, len(parent)))
        f.write("\n".join(lines) + ("\n" if trailing_newline else ""))


class ORBVocabulary:
    def __init__(self, k: int, L: int, parent, is_leaf, desc, weight, device: int = 0):
        self._L = _lib.lib()
        self._L.orb_vocabulary_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                  C.POINTER(C.c_void_p)]
        self._L.orb_vocabulary_destroy.argtypes = [C.c_void_p]
        self._L.orb_vocabulary_destroy.restype = None
        self._L.orb_vocabulary_words.argtypes = [C.c_void_p]
        self._L.orb_bow_transform.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 10
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, np.float64)
        assert len(parent) == len(is_leaf) == len(desc) == len(weight)
        self.k, self.L = k, L
        self._h = C.c_void_p()
        check(self._L.orb_vocabulary_create(k, L, len(parent), _vp(parent), _vp(is_leaf), _vp(desc), _vp(weight), device, C.byref(self._h)))

    @classmethod
    def loadFromTextFile(cls, path: str, device: int = 0) -> "ORBVocabulary":
        k, L, scoring, weighting, parent, is_leaf, desc, weight = read_text_vocabulary(path)
        if (scoring, weighting) != (0, 0):
            raise ValueError("only the ORB vocabulary configuration is supported: L1_NORM scoring (0), TF_IDF weighting (0)")
        return cls(k, L, parent, is_leaf, desc, weight, device)

    def __del__(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.orb_vocabulary_destroy(self._h)
            self._h = None

    def size(self) -> int:
        return self._L.orb_vocabulary_words(self._h)

    def transform(self, desc: np.ndarray, n: np.ndarray, levelsup: int = 4):
        """desc [F,cap,32] uint8, n [F]. Returns a dict of arrays in the C-ABI layout (word, node, bow_word, bow_value, bow_count,
        fv_node, fv_start, fv_feature, fv_count)."""
        desc = np.ascontiguousarray(desc, np.uint8)
        F, cap, _ = desc.shape
        n = np.ascontiguousarray(n, np.int32)
        r = dict(word=np.empty((F, cap), np.int32), node=np.empty((F, cap), np.int32), bow_word=np.empty((F, cap), np.int32),
                 bow_value=np.empty((F, cap), np.float64), bow_count=np.empty(F, np.int32), fv_node=np.empty((F, cap), np.int32),
                 fv_start=np.empty((F, cap + 1), np.int32), fv_feature=np.empty((F, cap), np.int32), fv_count=np.empty(F, np.int32))
        check(self._L.orb_bow_transform(self._h, MEM_HOST, _vp(desc), _vp(n), cap, F, levelsup, _vp(r["word"]), _vp(r["node"]),
                                        _vp(r["bow_word"]), _vp(r["bow_value"]), _vp(r["bow_count"]), _vp(r["fv_node"]),
                                        _vp(r["fv_start"]), _vp(r["fv_feature"]), _vp(r["fv_count"]), None))
        return r

    @staticmethod
    def frame_maps(r: dict, f: int):
        """(BowVector, FeatureVector) of frame f as the reference's std::map contents: {word: value}, {node: [features]}."""
        bc, fc = int(r["bow_count"][f]), int(r["fv_count"][f])
        bow = {int(w): float(v) for w, v in zip(r["bow_word"][f, :bc], r["bow_value"][f, :bc])}
        st = r["fv_start"][f]
        fv = {int(r["fv_node"][f, j]): r["fv_feature"][f, st[j]:st[j + 1]].tolist() for j in range(fc)}
        return bow, fv
