"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/gzb200.h
declares, its host-only arithmetic (bound, CRC combine, framing, corpus) agrees with the oracle, and it refuses to
compute without a GPU.  No GPU needed; no compute entry point is exercised here."""
import ctypes, os, re, zlib
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol(gz):
    hdr = open(os.path.join(ROOT, "include", "gzb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(gzb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 18
    L = ctypes.CDLL(gz.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # the input generator is a library of its own (include/gzb_corpus.h -> libgzbcorpus.so): the engine does not export it,
    # and a process that only generates input -- bench.py's reference arm -- maps nothing of the engine
    assert not hasattr(L, "gzb_corpus_fill") and not hasattr(L, "gzb_test_corrupt_at")
    assert hasattr(ctypes.CDLL(gz.CORPUS_LIB_PATH), "gzb_corpus_fill")


def test_header_cites_reference_for_each_entry_point():
    hdr = open(os.path.join(ROOT, "include", "gzb200.h")).read()
    assert hdr.count("src/parallel.c") >= 8 and "lib/zlib" in hdr and "src/zip.c:149-153" in hdr


def test_no_device_is_an_error_not_a_fallback(gz):
    if gz.lib().gzb_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(gz.GzbError) as e:
        gz.Context(0)
    assert e.value.code == gz.E_NO_DEVICE
    r = subprocess_run_cli(gz, b"hello")
    assert r.returncode != 0 and b"no usable CUDA device" in r.stderr and r.stdout == b""


def subprocess_run_cli(gz, data, *args):
    import subprocess
    return subprocess.run([os.path.join(os.path.dirname(gz.LIB_PATH), "gzb200"), *args], input=data, stdout=subprocess.PIPE, stderr=subprocess.PIPE)


def test_strerror_and_levels(gz):
    L = gz.lib()
    assert L.gzb_strerror(0) == b"ok"
    for code in range(-7, 0):
        assert L.gzb_strerror(code) not in (b"ok", b"unknown error")
    assert L.gzb_strerror(-99) == b"unknown error"


def test_bound_covers_all_stored_worst_case(gz):
    # worst case of the engine: every block stored (5 bytes per <=65535-byte sub-block) + one alignment block per chunk
    for n in (0, 1, 65535, 65536, 131072, 131073, 1 << 20, (4 << 30) + 17):
        chunks = max(1, -(-n // gz.CHUNK))
        assert gz.bound(n) >= n + chunks * (2 * 2 * 5 + 5)
        rchunks = max(1, -(-n // gz.RSYNC_CHUNK))
        assert gz.bound(n, gz.F_RSYNCABLE) >= n + rchunks * (5 + 5)
    # and zlib's own bound for its block sizes (lib/zlib/deflate.c:707-708)
    n = 1 << 26
    assert gz.bound(n) >= n + (n >> 12) // 4


@pytest.mark.parametrize("la,lb", [(0, 0), (3, 0), (0, 3), (131072, 131072), (131072, 5), (77, (1 << 22) + 1)])
def test_crc32_combine_matches_oracle(gz, orc, la, lb):
    rng = np.random.default_rng(la * 31 + lb)
    a = rng.integers(0, 256, la, dtype=np.uint8).tobytes(); b = rng.integers(0, 256, lb, dtype=np.uint8).tobytes()
    assert gz.crc32_combine(zlib.crc32(a), zlib.crc32(b), lb) == zlib.crc32(a + b) == orc.crc32_combine(zlib.crc32(a), zlib.crc32(b), lb)


def test_crc32_combine_huge_lengths(gz, orc):
    for n in ((1 << 32) + 5, (64 << 30) - 1, (1 << 40) + 12345):
        assert gz.crc32_combine(0xCAFEBABE, 0x1234, n) == orc.crc32_combine(0xCAFEBABE, 0x1234, n)


def test_framing_matches_oracle_and_golden(gz, orc, golden):
    for level in range(1, 10):
        assert gz.gzip_header(level) == orc.gzip_header(level)
        assert gz.gzip_header(level, 0x5F000000, b"a/b.txt") == orc.gzip_header(level, 0x5F000000, b"a/b.txt")
    assert gz.gzip_trailer(0xDEADBEEF, (5 << 32) + 9) == orc.gzip_trailer(0xDEADBEEF, (5 << 32) + 9)      # ISIZE mod 2^32
    ka = bytes.fromhex(golden["known_answers"]["parallel_nl_j4_L6"])
    assert gz.gzip_header(6) == ka[:10] and gz.gzip_trailer(zlib.crc32(b"parallel\n"), 9) == ka[-8:]


def test_corpus_is_deterministic_and_sliceable(gz, golden):
    import hashlib
    for kind, e in golden["corpora"].items():
        d = gz.corpus(kind, e["n"], seed=golden["seed"], threads=3)
        assert hashlib.sha256(d.tobytes()).hexdigest() == e["sha256"]
    whole = gz.corpus("mixed", 3 << 20, seed=9)
    part = gz.corpus("mixed", 1234567, seed=9, offset=777777, threads=2)
    assert (whole[777777:777777 + 1234567] == part).all()


def _plan(gz, n):
    L = ctypes.CDLL(gz.LIB_PATH)
    L.gzb_host_batch_plan.argtypes = [ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]
    L.gzb_host_batch_plan.restype = ctypes.c_int
    off = (ctypes.c_uint64 * 4096)()
    k = L.gzb_host_batch_plan(n, off, 4096)
    assert k >= 1
    return [int(off[i]) for i in range(k + 1)]


def test_host_batch_plan(gz, monkeypatch):
    """gzb_deflate_host's batch boundaries (host arithmetic): they tile [0, n) in order, every inner boundary is a multiple
    of both chunk sizes, no batch exceeds the full size, and a large input starts with growing batches (each at most 3x
    the one before, so its H2D copy hides behind the previous batch's kernels) and ends with a small one."""
    MB = 1 << 20
    for batch_mb, edge_mb in ((512, 64), (2, 1), (8, 3), (512, 0), (4, 64)):
        monkeypatch.setenv("GZB_HOST_BATCH_MB", str(batch_mb)); monkeypatch.setenv("GZB_HOST_EDGE_MB", str(edge_mb))
        B, E = batch_mb * MB, edge_mb * MB
        sizes = [0, 1, 131071, 131072, 131073, E, 4 * E, 4 * E + 1, 5 * E + 12345, B - 1, B, B + 1, 3 * B + 777, (4096 if batch_mb > 8 else 64) * MB,
                 (4096 if batch_mb > 8 else 64) * MB + 4321, 7 * B + 3 * E + 99]
        for n in sizes:
            off = _plan(gz, n)
            assert off[0] == 0 and off[-1] == n, (batch_mb, edge_mb, n)
            lens = [b - a for a, b in zip(off, off[1:])]
            assert all(l > 0 for l in lens) or n == 0
            assert all(o % (128 * 1024) == 0 and o % (16 * 1024) == 0 for o in off[:-1])
            assert max(lens) <= B
            if E and E < B and n > 4 * E:
                assert lens[0] == E and lens[-1] <= E
                for a, b in zip(lens, lens[1:-1]): assert b <= 3 * a            # the ramp and the full-size run
            else:
                assert all(l == B for l in lens[:-1])
    L = ctypes.CDLL(gz.LIB_PATH)
    L.gzb_host_batch_plan.argtypes = [ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]
    off = (ctypes.c_uint64 * 2)()
    monkeypatch.setenv("GZB_HOST_BATCH_MB", "1"); monkeypatch.setenv("GZB_HOST_EDGE_MB", "0")
    assert L.gzb_host_batch_plan(10 * MB, off, 2) == gz.E_ARG                   # needs 11 entries
