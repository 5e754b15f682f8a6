"""The engine's CPU model (oracle/engine_model.c -- the executable spec the CUDA kernels are compared with byte for
byte on the GPU box) must itself satisfy the parity contract: zlib decodes it, and its size stays within 3 % of the
reference at the same level (sizes frozen from the reference in tests/golden/)."""
import zlib
import pytest
from conftest import EDGE_SIZES


@pytest.mark.parametrize("kind", ["text", "zipf", "binary", "mixed", "random"])
def test_model_ratio_within_3pct_of_reference(gz, orc, golden, kind):
    e = golden["corpora"][kind]
    d = gz.corpus(kind, e["n"], seed=golden["seed"])
    for level in (1, 2, 3, 4, 5, 6, 7, 8, 9):
        raw = orc.engine_model(d, level)
        assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()
        ref_body = e["parallel"][str(level)]["size"] - 18
        assert raw.size <= 1.03 * ref_body, (kind, level, raw.size, ref_body)


@pytest.mark.parametrize("n", EDGE_SIZES)
def test_model_edge_sizes(gz, orc, n):
    d = gz.corpus("text", n, seed=3)
    for level in (1, 6):
        raw = orc.engine_model(d, level)
        assert orc.inflate(raw, n).tobytes() == d.tobytes()
        assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()


def test_model_fragments_are_byte_aligned_and_independent_of_the_cut(gz, orc):
    # chunk grid semantics of parallel.c: compressing [0,n) in one call equals compressing two shards with a 32 KiB halo
    n = 5 * 131072 + 4321
    d = gz.corpus("zipf", n, seed=11)
    whole = orc.engine_model(d, 6).tobytes()
    cut = 3 * 131072
    a = orc.engine_model(d[:cut].copy(), 6, final=False).tobytes()
    b = orc.engine_model(d, 6, halo=32768, skip=cut).tobytes()
    assert a + b == whole
    assert zlib.decompress(a + b, -15) == d.tobytes()


def test_model_rsyncable(gz, orc, golden):
    e = golden["corpora"]["text"]
    d = gz.corpus("text", e["n"], seed=golden["seed"])
    raw = orc.engine_model(d, 6, rsyncable=True)
    assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()
    assert raw.size <= 1.03 * (e["rsyncable_parallel_6"] - 18)


def test_model_static_blocks_with_high_literals(gz, orc):
    """Static blocks carrying literals 144..255 (9-bit codes, which depend on symbols 286/287 being part of the code,
    RFC 1951 3.2.6): tiny binary inputs and a tiny last chunk after incompressible ones."""
    import numpy as np
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 5, 17, 40):
        for _ in range(8):
            d = rng.integers(144, 256, n, dtype=np.uint8)
            for level in (1, 6, 9):
                raw = orc.engine_model(d, level)
                assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()
                assert orc.inflate(raw, n).tobytes() == d.tobytes()
    for n in (2 * 131072 + 1, 2 * 131072 + 5, 131072 + 3):
        d = gz.corpus("random", n, seed=9)
        raw = orc.engine_model(d, 6)
        assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()


def test_model_stored_decision(gz, orc):
    """gzm_chunk_is_stored (the CPU statement of k_scan's test): random and already-compressed data are stored without a
    search; text, base64 and flat-but-repetitive data are not."""
    import base64, ctypes
    import numpy as np
    L = orc.lib(); L.gzm_chunk_is_stored.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    f = lambda d, D=0: L.gzm_chunk_is_stored(d.ctypes.data, D, d.size - D)
    rnd = gz.corpus("random", 131072 + 32768, seed=1)
    assert f(rnd, 32768) == 1 and f(rnd[:131072].copy()) == 1
    assert f(gz.corpus("text", 131072)) == 0 and f(gz.corpus("binary", 131072)) == 0 and f(gz.corpus("zipf", 131072)) == 0
    z = np.frombuffer(zlib.compress(gz.corpus("text", 1 << 20).tobytes(), 9)[:131072], dtype=np.uint8).copy()
    assert f(z) == 1
    assert f(np.concatenate([rnd[:16384]] * 8)) == 0                                   # the same 16 KiB eight times
    assert f(np.concatenate([rnd[:32768], rnd[:32768], gz.corpus("random", 98304, seed=7)]), 32768) == 0   # starts with a copy of its dictionary
    assert f(np.frombuffer(base64.b64encode(rnd[:98304].tobytes()), dtype=np.uint8).copy()) == 0
    assert f(rnd[:4095].copy()) == 0                                                   # below SCAN_MIN: always the normal path
