"""Runs oracle/_ref/orb_ref_extract: the reference's own src/ORBextractor.cc compiled unmodified against oracle/cvshim
(built by `make -C oracle _ref` where /root/reference exists; the binary is git-ignored and travels with the snapshot).
Test infrastructure only."""
import os
import struct
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_extract")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
                     ("class_id", "<i4")])


def available() -> bool:
    return os.access(BIN, os.X_OK)


def ref_extract(frames: np.ndarray, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, legacy_blur=False):
    """frames: [n, H, W] u8 -> (tables [4, nlevels] f32, list of (kps, desc, [padded level images]))."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, H, W = frames.shape
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<8if", 0x4F524231, n, W, H, nfeatures, nlevels, ini_th, min_th, scale_factor))
            f.write(frames.tobytes())
        env = dict(os.environ)
        env.pop("ORB_REF_LEGACY_BLUR", None)
        if legacy_blur:
            env["ORB_REF_LEGACY_BLUR"] = "1"
        subprocess.run([BIN, fin, fout], check=True, env=env, timeout=600)
        buf = open(fout, "rb").read()
    pos = 0
    nl = struct.unpack_from("<i", buf, pos)[0]; pos += 4
    tables = np.frombuffer(buf, "<f4", 4 * nl, pos).reshape(4, nl).copy(); pos += 16 * nl
    out = []
    for _ in range(n):
        N = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        kps = np.frombuffer(buf, KP_DTYPE, N, pos).copy(); pos += 28 * N
        desc = np.frombuffer(buf, np.uint8, 32 * N, pos).reshape(N, 32).copy(); pos += 32 * N
        levels = []
        for _l in range(nl):
            c, r = struct.unpack_from("<2i", buf, pos); pos += 8
            levels.append(np.frombuffer(buf, np.uint8, r * c, pos).reshape(r, c).copy()); pos += r * c
        out.append((kps, desc, levels))
    assert pos == len(buf)
    return tables, out


def write_input(path, frames, nfeatures, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, H, W = frames.shape
    with open(path, "wb") as f:
        f.write(struct.pack("<8if", 0x4F524231, n, W, H, nfeatures, nlevels, ini_th, min_th, scale_factor))
        f.write(frames.tobytes())


BOW_BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_bow")


def bow_available() -> bool:
    return os.access(BOW_BIN, os.X_OK)


def ref_bow(voc_path: str, frames, levelsup: int = 4):
    """The reference's own DBoW2: loadFromTextFile(voc_path) + transform(). frames: list of [n, 32] u8 arrays ->
    list of (BowVector {word: value}, FeatureVector {node: [feature indices]})."""
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "desc.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<2i", len(frames), levelsup))
            for a in frames:
                a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
                f.write(struct.pack("<i", len(a)))
                f.write(a.tobytes())
        subprocess.run([BOW_BIN, voc_path, fin, fout], check=True, timeout=600)
        buf = open(fout, "rb").read()
    pos, out = 0, []
    for _ in frames:
        nb = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        bow = {}
        for _j in range(nb):
            w, v = struct.unpack_from("<Id", buf, pos); pos += 12
            bow[w] = v
        nf = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        fv = {}
        for _j in range(nf):
            node, m = struct.unpack_from("<Ii", buf, pos); pos += 8
            fv[node] = list(struct.unpack_from(f"<{m}I", buf, pos)); pos += 4 * m
        out.append((bow, fv))
    assert pos == len(buf)
    return out
