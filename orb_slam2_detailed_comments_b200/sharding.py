"""How the two paths partition over the GPUs of one box.

Extraction: frames are independent -> contiguous frame shards, no data-path collective.
kNN: database rows sharded contiguously and in ascending rank order (so "earliest global index wins ties" survives
the merge); queries replicated; one all_gather of the per-query top-2 records, then a local merge.
Pure index arithmetic: no compute happens here."""
from __future__ import annotations


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """[begin, end) of rank's contiguous shard; the first `total % world` ranks get one extra unit."""
    if world < 1 or not 0 <= rank < world or total < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def frame_shard(n_frames: int, rank: int, world: int) -> tuple[int, int]:
    """Frames (or stereo pairs, when n_frames counts pairs) of this rank."""
    return shard_range(n_frames, rank, world)


def db_shard(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Database rows of this rank; `begin` is the index_base passed to orb_hamming_knn."""
    return shard_range(n_rows, rank, world)
