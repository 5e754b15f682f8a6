"""Multi-GPU sharding of the chunk grid (host logic).

The path partitions by independent chunks with zero exchange steps: rank r of `world` takes a contiguous range of
128 KiB chunks, reads 32 KiB of halo before it (the dictionary of its first chunk -- what the reference memcpy's from
the previous read, src/parallel.c:767-774), and the host concatenates the fragment runs in rank order and folds the
CRCs with crc32_combine (src/parallel.c:544).  No collective is involved."""
from . import CHUNK, DICT


def shard_plan(n: int, rank: int, world: int, chunk: int = CHUNK, dict_size: int = DICT):
    """-> (lo, hi, halo): byte range [lo, hi) of rank `rank` and the number of halo bytes before lo."""
    nchunks = -(-n // chunk)
    per, extra = divmod(nchunks, world)
    c0 = rank * per + min(rank, extra)
    c1 = c0 + per + (1 if rank < extra else 0)
    lo, hi = min(c0 * chunk, n), min(c1 * chunk, n)
    return lo, hi, (dict_size if lo > 0 else 0)
