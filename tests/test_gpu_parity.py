"""Parity tests proper (B200 only): the CUDA path, called through the C ABI, against the oracle.
Bars (BASELINE.json north_star): every output decodes bit-exactly with the reference gzip -d, zlib and the oracle's
own decoder; CRC32/ISIZE bit-exact; size <= 1.03 x the reference at the same level; and -- stronger than the
reference's own tests -- the bytes equal the engine's CPU model, so nothing about the result depends on GPU scheduling."""
import os, subprocess, zlib
import numpy as np
import pytest
from conftest import EDGE_SIZES

pytestmark = pytest.mark.gpu
RATIO_TOL = 1.03          # BASELINE.json: "<=3% larger than the reference at the same level"


def check_member(gz, orc, data, raw, r, level, flags_final=True):
    n = data.size
    assert r.in_bytes == n and r.out_bytes == raw.size
    assert r.crc32 == zlib.crc32(data.tobytes()) == orc.crc32(data)
    if flags_final:
        assert zlib.decompress(raw.tobytes(), -15) == data.tobytes()                 # zlib
        back = orc.inflate(raw, n)
        assert back is not None and back.tobytes() == data.tobytes()                 # oracle decoder
        member = gz.gzip_header(level) + raw.tobytes() + gz.gzip_trailer(r.crc32, n)
        assert member[-8:] == orc.gzip_trailer(orc.crc32(data), n)                   # CRC32 / ISIZE byte-exact
        if orc.have_ref():
            assert orc.ref_gunzip(member) == data.tobytes()                          # the reference's own gzip -d
        p = subprocess.run(["gunzip", "-c"], input=member, stdout=subprocess.PIPE)   # system gunzip
        assert p.returncode == 0 and p.stdout == data.tobytes()


@pytest.mark.parametrize("n", EDGE_SIZES)
def test_edge_sizes(gz, orc, ctx, n):
    d = gz.corpus("text", n, seed=3)
    for level in (1, 6, 9):
        raw, r = ctx.deflate_host(d, level)
        check_member(gz, orc, d, raw, r, level)
        assert raw.tobytes() == orc.engine_model(d, level).tobytes()


@pytest.mark.parametrize("kind", ["text", "zipf", "binary", "mixed", "random"])
def test_golden_corpora_all_levels(gz, orc, ctx, golden, kind):
    e = golden["corpora"][kind]
    d = gz.corpus(kind, e["n"], seed=golden["seed"])
    for level in range(1, 10):
        raw, r = ctx.deflate_host(d, level)
        check_member(gz, orc, d, raw, r, level)
        assert r.crc32 == e["crc32"]
        ref_body = e["parallel"][str(level)]["size"] - 18
        assert raw.size <= RATIO_TOL * ref_body, (kind, level, raw.size, ref_body)
        assert raw.tobytes() == orc.engine_model(d, level).tobytes()
        assert gz.gzip_header(level).hex() == e["parallel"][str(level)]["head"]
        assert gz.gzip_trailer(r.crc32, r.in_bytes).hex() == e["parallel"][str(level)]["tail"]


def test_known_answers(gz, orc, ctx, golden):
    # empty input: the valid 20-byte member of the single-stream path (the reference's -j path deadlocks)
    raw, r = ctx.deflate_host(np.empty(0, dtype=np.uint8), 6)
    assert (gz.gzip_header(6) + raw.tobytes() + gz.gzip_trailer(r.crc32, 0)).hex() == golden["known_answers"]["empty_single_L6"]
    # tests/parallel's 9-byte case and tests/memcpy-abuse's window-edge case
    for data in (b"parallel\n", b"wxy" + b"%032767d" % 0):
        d = np.frombuffer(data, dtype=np.uint8)
        raw, r = ctx.deflate_host(d, 6)
        check_member(gz, orc, d, raw, r, 6)
    raw, r = ctx.deflate_host(np.frombuffer(b"wxy" + b"%032767d" % 0, dtype=np.uint8), 6)
    assert raw.size + 18 <= golden["known_answers"]["memcpy_abuse_L6_j2_size"] + 8


def test_degenerate_inputs(gz, orc, ctx):
    for d in (np.zeros(400000, dtype=np.uint8), np.full(131072 + 77, 0x41, dtype=np.uint8),
              np.tile(np.arange(256, dtype=np.uint8), 3000), np.tile(np.frombuffer(b"abc", dtype=np.uint8), 100001)):
        for level in (1, 6, 9):
            raw, r = ctx.deflate_host(d, level)
            check_member(gz, orc, d, raw, r, level)
            assert raw.tobytes() == orc.engine_model(d, level).tobytes()


def _lz_structured(rng, n, max_dist, lit_run=12):
    """bytes built the way an LZ77 decoder builds them: short literal runs and copies at chosen distances and lengths"""
    out = bytearray(rng.integers(0, 256, 64, dtype=np.uint8).tobytes())
    while len(out) < n:
        if rng.random() < 0.35:
            out += rng.integers(97, 123, int(rng.integers(1, lit_run)), dtype=np.uint8).tobytes()
        else:
            dist = int(min(len(out), rng.choice([1, 2, 3, 4, 5, 6, 7, 255, 256, 257, 258, 259, 4096, 32766, 32767, 32768, int(rng.integers(1, max_dist))])))
            ln = int(rng.choice([3, 4, 5, 6, 7, 8, 16, 17, 257, 258, 259, 300, int(rng.integers(3, 600))]))
            for _ in range(ln):
                out.append(out[-dist])
    return np.frombuffer(bytes(out[:n]), dtype=np.uint8).copy()


@pytest.mark.parametrize("seed", range(6))
def test_structured_inputs_equal_the_model(gz, orc, ctx, seed):
    """Inputs aimed at the edges of the match finder: periods around the longest match (258) and the window (32 KiB),
    copies that straddle tile (2 KiB), block (64 KiB) and chunk (128 KiB) boundaries, distances of exactly 32768 and one
    beyond, and chunks of random bytes between compressible ones (k_lz77's segments).  Bytes must equal the sequential
    model and decode with zlib."""
    rng = np.random.default_rng(1000 + seed)
    cases = []
    for period in (1, 2, 3, 5, 6, 7, 257, 258, 259, 2047, 2048, 2049, 32767, 32768, 32769, 65536)[seed::6]:
        unit = rng.integers(0, 256, period, dtype=np.uint8)
        cases.append((f"period {period}", np.tile(unit, (3 * gz.CHUNK + 4097) // period + 1)[: 3 * gz.CHUNK + 4097]))
    cases.append(("lz near", _lz_structured(rng, 2 * gz.CHUNK + int(rng.integers(1, 5000)), 300)))
    cases.append(("lz far", _lz_structured(rng, 3 * gz.CHUNK + int(rng.integers(1, 5000)), 40000)))
    parts = []                                                   # text / random / text / random ... in chunk-sized and ragged pieces
    for k in range(7):
        ln = gz.CHUNK if k % 3 else int(rng.integers(1000, 2 * gz.CHUNK))
        parts.append(rng.integers(0, 256, ln, dtype=np.uint8) if (k + seed) % 2 else gz.corpus("text", ln, seed=seed * 16 + k))
    cases.append(("segments", np.concatenate(parts)))
    for name, d in cases:
        for level in (1, 5, 9):
            raw, r = ctx.deflate_host(d, level)
            assert r.crc32 == zlib.crc32(d.tobytes())
            assert raw.tobytes() == orc.engine_model(d, level).tobytes(), (name, level, d.size)
            assert zlib.decompress(raw.tobytes(), -15) == d.tobytes(), (name, level)


def test_rsyncable(gz, orc, ctx, golden):
    e = golden["corpora"]["text"]
    d = gz.corpus("text", e["n"], seed=golden["seed"])
    raw, r = ctx.deflate_host(d, 6, gz.F_FINAL | gz.F_RSYNCABLE)
    check_member(gz, orc, d, raw, r, 6)
    assert raw.tobytes() == orc.engine_model(d, 6, rsyncable=True).tobytes()
    assert raw.size <= RATIO_TOL * (e["rsyncable_parallel_6"] - 18)


def test_shards_with_halo_concatenate_to_the_whole(gz, orc, ctx):
    """Multi-GPU contract: contiguous shards, 32 KiB halo, host concatenation + crc32_combine."""
    import torch
    n = 9 * gz.CHUNK + 1000
    d = gz.corpus("mixed", n, seed=21)
    whole, rw = ctx.deflate_host(d, 6)
    t = torch.from_numpy(d).cuda()
    cuts = [0, 2 * gz.CHUNK, 5 * gz.CHUNK, n]
    parts, crc = [], 0
    for i in range(3):
        lo, hi = cuts[i], cuts[i + 1]
        out = torch.empty(gz.bound(hi - lo), dtype=torch.uint8, device="cuda")
        r = ctx.deflate_device(t[lo:], hi - lo, out, 6, flags=gz.F_FINAL if i == 2 else 0, halo=gz.DICT if lo else 0)
        parts.append(out[: r.out_bytes].cpu().numpy().tobytes())
        crc = gz.crc32_combine(crc, r.crc32, hi - lo)
    assert b"".join(parts) == whole.tobytes()
    assert crc == rw.crc32 == zlib.crc32(d.tobytes())


def test_device_entry_points(gz, orc, ctx):
    import torch
    n = 3 * gz.CHUNK + 5
    d = gz.corpus("zipf", n, seed=2)
    t = torch.from_numpy(d).cuda()
    assert ctx.crc32_device(t, n) == zlib.crc32(d.tobytes())
    assert ctx.crc32_host(d) == zlib.crc32(d.tobytes())
    assert ctx.crc32_host(d[1000:].copy(), ctx.crc32_host(d[:1000].copy())) == zlib.crc32(d.tobytes())
    out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ctx.deflate_device_async(t, n, out, 6, stream=s.cuda_stream)
    r = ctx.wait()
    assert out[: r.out_bytes].cpu().numpy().tobytes() == orc.engine_model(d, 6).tobytes()
    ms = ctx.stage_ms()
    assert ms["lz77"] > 0 and ms["total"] >= ms["lz77"]


def test_error_codes(gz, ctx):
    import torch
    t = torch.zeros(4096, dtype=torch.uint8, device="cuda"); small = torch.empty(16, dtype=torch.uint8, device="cuda")
    big = torch.empty(gz.bound(4096), dtype=torch.uint8, device="cuda")
    for level in (0, 10, -1):
        with pytest.raises(gz.GzbError) as e:
            ctx.deflate_device(t, 4096, big, level)
        assert e.value.code == gz.E_LEVEL
    with pytest.raises(gz.GzbError) as e:
        ctx.deflate_device(t, 4096, small, 6)
    assert e.value.code == gz.E_OUT_FULL
    with pytest.raises(gz.GzbError) as e:
        ctx.deflate_device(t[1:], 4095, big, 6)             # misaligned input
    assert e.value.code == gz.E_ARG
    with pytest.raises(gz.GzbError) as e:
        ctx.deflate_device(t, 4096, big, 6, halo=100)
    assert e.value.code == gz.E_ARG


def test_incompressible_chunks_are_stored_without_a_search(gz, orc, ctx, monkeypatch):
    """k_scan decides "stored" per chunk BEFORE the match finder (flat byte histogram and no repeated 8-byte grams);
    the model takes the same decision from the same integers.  Flat but repetitive data must still be searched."""
    rnd = gz.corpus("random", 5 * gz.CHUNK + 4321, seed=33)
    rep = np.concatenate([rnd[:16384]] * 24 + [rnd[:777]])                          # random, but every 16 KiB the same
    z = np.frombuffer(zlib.compress(gz.corpus("text", 3 << 20, seed=34).tobytes(), 9), dtype=np.uint8).copy()   # deflate output: flat, no repeats
    mix = np.concatenate([gz.corpus("text", 2 * gz.CHUNK, seed=35), rnd[: 2 * gz.CHUNK], gz.corpus("binary", gz.CHUNK + 5, seed=36), rnd[: gz.CHUNK + 1]])
    for d in (rnd, rep, z, mix):
        for level in (1, 6, 9):
            raw, r = ctx.deflate_host(d, level, gz.F_FINAL | gz.F_VERIFY)
            check_member(gz, orc, d, raw, r, level)
            assert raw.tobytes() == orc.engine_model(d, level).tobytes()
    raw, r = ctx.deflate_host(rnd, 6)
    assert raw.size <= rnd.size + 20 * (rnd.size // gz.CHUNK + 1)                      # all stored
    raw, r = ctx.deflate_host(rep, 6)
    assert raw.size < rep.size // 10                                                   # searched: the repeats are found
    # with the test switched off every chunk takes the normal path: same data back, size within a few bytes per block
    monkeypatch.setenv("GZB_SCAN", "0")
    c = gz.Context(0)
    try:
        for d in (rnd, mix):
            raw0, r0 = c.deflate_host(d, 6, gz.F_FINAL | gz.F_VERIFY)
            assert zlib.decompress(raw0.tobytes(), -15) == d.tobytes()
            raw1, _ = ctx.deflate_host(d, 6)
            assert abs(int(raw0.size) - int(raw1.size)) <= 64 * (d.size // gz.CHUNK + 1)
    finally:
        c.close()


def test_host_call_over_several_contexts(gz, orc):
    """gzb_deflate_host_multi: contiguous shards on chunk boundaries, one context (GPU) each, 32 KiB halo, host
    crc32_combine; compacted, the bodies are exactly the one-call body.  (One GPU is enough: contexts may share a device.)"""
    import torch
    ndev = torch.cuda.device_count()
    for nctx in (1, 2, 3, 5):
        cs = [gz.Context(i % ndev) for i in range(nctx)]
        try:
            for n in (0, 1, gz.CHUNK, 2 * gz.CHUNK + 1, 11 * gz.CHUNK + 4321):
                d = gz.corpus("mixed", n, seed=51)
                want = orc.engine_model(d, 6).tobytes() if n else None
                raw, r, sh = gz.deflate_host_multi(cs, d, 6)
                assert r.in_bytes == n and r.crc32 == zlib.crc32(d.tobytes())
                if n:
                    assert raw.tobytes() == want
                else:
                    assert zlib.decompress(raw.tobytes(), -15) == b""
                assert sum(x[1] for x in sh) == n and all(x[0] % gz.CHUNK == 0 for x in sh if x[1])
                # not compacted: every body where the shard table says
                out, r2, sh2 = gz.deflate_host_multi(cs, d, 6, compact=False)
                assert b"".join(out[o: o + l].tobytes() for _, _, o, l in sh2) == raw.tobytes()
        finally:
            for c in cs: c.close()


def test_host_call_with_a_smaller_output_buffer(gz, orc, ctx):
    """gzb_deflate_host accepts out_cap < gzb_bound(n) (a caller that knows its ratio need not page-lock n bytes of output)
    and reports GZB_E_OUT_FULL instead of writing past it; the context stays usable."""
    d = gz.corpus("text", 3 * gz.CHUNK + 7, seed=12)
    want = orc.engine_model(d, 6).tobytes()
    out = np.empty(len(want) + 100, dtype=np.uint8)
    raw, r = ctx.deflate_host(d, 6, out=out)
    assert raw.tobytes() == want
    guard = np.full(4096, 0x5A, dtype=np.uint8)
    with pytest.raises(gz.GzbError) as e:
        ctx.deflate_host(d, 6, out=guard[:1000])
    assert e.value.code == gz.E_OUT_FULL and (guard[1000:] == 0x5A).all()
    raw, r = ctx.deflate_host(d, 6, out=out)
    assert raw.tobytes() == want


def test_host_driver_and_cli(gz, orc, tmp_path):
    """gzb_zip_fd / the gzb200 CLI, in the style of the reference's tests/parallel: compress | gzip -d | cmp."""
    exe = os.path.join(os.path.dirname(gz.LIB_PATH), "gzb200")
    for data in (b"", b"parallel\n", gz.corpus("text", 700001, seed=4).tobytes()):
        for args in (["-6"], ["-1"], ["-9"], ["-6", "--rsyncable"]):
            p = subprocess.run([exe, *args], input=data, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            assert p.returncode == 0, p.stderr
            assert subprocess.run(["gunzip", "-c"], input=p.stdout, stdout=subprocess.PIPE, check=True).stdout == data
            if orc.have_ref():
                assert orc.ref_gunzip(p.stdout) == data
    src = tmp_path / "in.bin"; dst = tmp_path / "out.gz"
    payload = gz.corpus("mixed", (2 << 20) + 3, seed=8).tobytes(); src.write_bytes(payload)
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6, name=b"in.bin", mtime=1234567)
    m = dst.read_bytes()
    assert bi == len(payload) and bo == len(m)
    assert m[:10] == orc.gzip_header(6, 1234567, b"in.bin")[:10] and m[10:17] == b"in.bin\0"
    assert zlib.decompress(m, 31) == payload


def test_large_roundtrip_properties(gz, orc, ctx):
    """Full-size behaviour through size-independent properties: round trip, CRC of the whole = combine of shard CRCs,
    per-call determinism."""
    import torch
    n = 1 << 30
    d = gz.corpus("text", n, seed=77)
    t = torch.from_numpy(d).cuda()
    out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    r = ctx.deflate_device(t, n, out, 6)
    raw = out[: r.out_bytes].cpu().numpy()
    assert r.crc32 == zlib.crc32(d.tobytes())
    assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()
    r2 = ctx.deflate_device(t, n, out, 6)
    assert r2.out_bytes == r.out_bytes and (out[: r.out_bytes].cpu().numpy() == raw).all()
    ref = orc.ref_gzip(d[: 64 << 20], 6, threads=os.cpu_count() or 8) if orc.have_ref() else None
    if ref is not None:
        mine, _ = ctx.deflate_host(d[: 64 << 20].copy(), 6)
        assert mine.size <= RATIO_TOL * (len(ref) - 18)


def test_host_pipeline_batches_and_sub_batches(gz, orc, monkeypatch):
    """The host path streams batches through two device slots, and a device call is cut into sub-batches chained by
    a device-side cursor: with both made tiny, the bytes must still equal the one-shot result (and the model).
    (The knobs are read once, by gzb_create: every setting gets its own context.)"""
    n = 23 * gz.CHUNK + 4321
    d = gz.corpus("mixed", n, seed=31)
    want = orc.engine_model(d, 6).tobytes()
    want_rs = orc.engine_model(d, 6, rsyncable=True).tobytes()
    monkeypatch.setenv("GZB_BATCH_CHUNKS", "3")
    for mb in ("1", "2", "512"):
        monkeypatch.setenv("GZB_HOST_BATCH_MB", mb)
        c = gz.Context(0)
        try:
            raw, r = c.deflate_host(d, 6)
            assert raw.tobytes() == want and r.crc32 == zlib.crc32(d.tobytes()) and r.in_bytes == n
            raw, r = c.deflate_host(d, 6, gz.F_FINAL | gz.F_RSYNCABLE)
            assert raw.tobytes() == want_rs and r.crc32 == zlib.crc32(d.tobytes())
        finally:
            c.close()
    # a large input starts and ends with a small batch (GZB_HOST_EDGE_MB) around the full-size ones
    n2 = 45 * gz.CHUNK + 777
    d2 = gz.corpus("text", n2, seed=32)
    want2 = orc.engine_model(d2, 6).tobytes()
    monkeypatch.setenv("GZB_HOST_BATCH_MB", "2"); monkeypatch.setenv("GZB_HOST_EDGE_MB", "1")
    c = gz.Context(0)
    try:
        raw, r = c.deflate_host(d2, 6)
        assert raw.tobytes() == want2 and r.crc32 == zlib.crc32(d2.tobytes()) and r.in_bytes == n2
    finally:
        c.close()


@pytest.mark.parametrize("pw,help_", [("8", "0"), ("8", "1"), ("8", "2"), ("8", "3"), ("16", "0"), ("16", "3")])
def test_kernel_variants_give_the_same_bytes(gz, orc, monkeypatch, pw, help_):
    """k_lz77 is built with a parser group of 8 or 16 warps, and the inserter / parser warps may or may not take search
    groups when their own tile is done (GZB_HELP bits); chosen by level, forced here at every level: none may change a byte."""
    monkeypatch.setenv("GZB_PW", pw); monkeypatch.setenv("GZB_HELP", help_)
    ctx = gz.Context(0)
    try:
        for kind, n in (("text", 5 * gz.CHUNK + 777), ("binary", 2 * gz.CHUNK), ("random", gz.CHUNK + 1)):
            d = gz.corpus(kind, n, seed=5)
            for level in (1, 4, 6, 9):
                raw, r = ctx.deflate_host(d, level)
                assert raw.tobytes() == orc.engine_model(d, level).tobytes(), (kind, level)
        # runs of chained chunks at full occupancy (the launch picks the run length: 5 here)
        d = gz.corpus("text", 700 * gz.CHUNK + 99, seed=6)
        raw, r = ctx.deflate_host(d, 6)
        assert r.crc32 == zlib.crc32(d.tobytes()) and zlib.decompress(raw.tobytes(), -15) == d.tobytes()
        want = orc.engine_model(d, 6).tobytes()
        assert raw.tobytes() == want
    finally:
        ctx.close()
    # the longest runs a launch ever picks (32 chunks per CTA, what a 4 GiB call uses), and odd ones
    for run in ("32", "13", "1"):
        monkeypatch.setenv("GZB_FORCE_RUN", run)
        ctx = gz.Context(0)
        try:
            raw, r = ctx.deflate_host(d, 6)
            assert raw.tobytes() == want, run
        finally:
            ctx.close()


@pytest.mark.parametrize("layout", ["0", "1", "2", "3"])
def test_warp_role_layouts_give_the_same_bytes(gz, orc, monkeypatch, layout):
    """What each warp of a k_lz77 CTA does (inserters, parsers, search; which scheduler they sit on) is a launch parameter
    (lz77_role_map, GZB_ROLEMAP); scheduling only, never a byte."""
    monkeypatch.setenv("GZB_ROLEMAP", layout)
    ctx = gz.Context(0)
    try:
        for kind, n in (("text", 5 * gz.CHUNK + 777), ("mixed", 3 * gz.CHUNK + 1)):
            d = gz.corpus(kind, n, seed=7)
            for level in (1, 4, 6, 9):
                raw, r = ctx.deflate_host(d, level)
                assert raw.tobytes() == orc.engine_model(d, level).tobytes(), (kind, level)
    finally:
        ctx.close()


def test_zip_fd_pipeline(gz, orc, tmp_path, monkeypatch):
    """gzb_zip_fd above its small-input threshold: reader / worker contexts / in-order writer over many batches, from a
    regular file (parallel pread) and from a pipe; the member must be header + the one-shot body + trailer."""
    import threading
    n = (37 << 20) + 12345
    d = gz.corpus("mixed", n, seed=41)
    want = orc.engine_model(d, 6).tobytes()
    src = tmp_path / "big.bin"; src.write_bytes(d.tobytes())
    # file to file takes the path without a serial stage (every worker reads its own batches and writes its fragments at
    # their offsets); GZB_ZIP_NO_DIRECT=1 keeps the one-reader / in-order-writer pipeline that pipes use
    for mb, serial in (("1", False), ("3", False), ("128", False), ("1", True), ("5", True)):
        monkeypatch.setenv("GZB_ZIP_BATCH_MB", mb)
        if serial:
            monkeypatch.setenv("GZB_ZIP_NO_DIRECT", "1")
        else:
            monkeypatch.delenv("GZB_ZIP_NO_DIRECT", raising=False)
        dst = tmp_path / f"out{mb}.gz"
        with open(src, "rb") as fi, open(dst, "wb") as fo:
            bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6)
            assert fi.tell() == n                       # the descriptor is left at EOF, as after a sequential read
        m = dst.read_bytes()
        assert bi == n and bo == len(m)
        assert m[:10] == gz.gzip_header(6) and m[10:-8] == want and m[-8:] == gz.gzip_trailer(zlib.crc32(d.tobytes()), n)
    monkeypatch.delenv("GZB_ZIP_NO_DIRECT", raising=False)
    # the contexts and pinned slots of the last call are still in the pool: the same call again, then an explicit release
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6)
    assert dst.read_bytes() == m
    gz.zip_release()
    with open(src, "rb") as fi, open(dst, "r+b") as fo:         # an output file that already holds bytes: written from the current offset
        fo.seek(7)
        bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6)
        assert fo.tell() == 7 + len(m)
    assert dst.read_bytes()[7:] == m
    # a pipe on both sides
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "2")
    r_in, w_in = os.pipe(); r_out, w_out = os.pipe()
    got = bytearray()
    def feed():
        with os.fdopen(w_in, "wb") as f: f.write(d.tobytes())
    def drain():
        with os.fdopen(r_out, "rb") as f: got.extend(f.read())
    t1 = threading.Thread(target=feed); t2 = threading.Thread(target=drain); t1.start(); t2.start()
    bi, bo = gz.zip_fd(r_in, w_out, level=6)
    os.close(w_out); os.close(r_in); t1.join(); t2.join()
    assert bi == n and bytes(got[10:-8]) == want and zlib.decompress(bytes(got), 31) == d.tobytes()
    # errors are reported: unwritable output
    with open(src, "rb") as fi, open(tmp_path / "ro", "wb") as fo:
        pass
    with open(src, "rb") as fi, open(tmp_path / "ro", "rb") as fo:
        with pytest.raises(gz.GzbError) as e:
            gz.zip_fd(fi.fileno(), fo.fileno(), level=6)
        assert e.value.code == gz.E_IO


def test_zip_fd_file_to_file_variants(gz, orc, ctx, tmp_path, monkeypatch):
    """The file-to-file path of gzb_zip_fd away from its defaults: --rsyncable (16 KiB chunks, no dictionary between
    batches), FNAME / MTIME in the header, an input descriptor that does not stand at the start of its file, more workers
    than batches, several workers per GPU, a single I/O thread, fragments written with pwrite instead of through a mapping.  Always: header + the one-shot body + trailer."""
    n = (41 << 20) + 4321
    d = gz.corpus("mixed", n, seed=77)
    src = tmp_path / "in.bin"; src.write_bytes(b"skip me" * 1000 + d.tobytes())
    skip = 7000
    def run(**kw):
        dst = tmp_path / "o.gz"
        with open(src, "rb") as fi, open(dst, "wb") as fo:
            fi.seek(skip)
            bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), **kw)
            assert fi.tell() == skip + n
        m = dst.read_bytes()
        assert bi == n and bo == len(m)
        return m
    crc = zlib.crc32(d.tobytes())
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "4")
    body6 = orc.engine_model(d, 6).tobytes()
    m = run(level=6, name=b"in.bin", mtime=0x01020304)
    assert m[:10] == bytes([31, 139, 8, 8, 4, 3, 2, 1, 0, 3]) and m[10:17] == b"in.bin\0"
    assert m[17:-8] == body6 and m[-8:] == gz.gzip_trailer(crc, n)
    for env in ({"GZB_ZIP_WORKERS_PER_GPU": "3"}, {"GZB_ZIP_IO_THREADS": "1"}, {"GZB_ZIP_BATCH_MB": "64"}, {"GZB_ZIP_WRITERS": "3", "GZB_ZIP_IO_PIECE_KB": "64", "GZB_ZIP_NO_MMAP": "1"}, {"GZB_ZIP_NO_MMAP": "1"}):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        m = run(level=6)
        assert m[10:-8] == body6 and m[-8:] == gz.gzip_trailer(crc, n), env
        for k in env:
            monkeypatch.delenv(k)
        monkeypatch.setenv("GZB_ZIP_BATCH_MB", "4")
    for ng in (2, 8):                                       # clamped to the GPUs present; on a multi-GPU box the batches alternate between them
        m = run(level=6, n_gpus=ng)
        assert m[10:-8] == body6 and m[-8:] == gz.gzip_trailer(crc, n), ng
    raw, r = ctx.deflate_host(d, 6, gz.F_FINAL | gz.F_RSYNCABLE)
    m = run(level=6, rsyncable=True)
    assert m[10:-8] == raw.tobytes() and zlib.decompress(m, 31) == d.tobytes()
    gz.zip_release()


def test_zip_fd_member_per_batch(gz, orc, tmp_path, monkeypatch):
    """opts.members (BASELINE config 5's "per-GPU members concatenated"): every batch a complete member.
    Multi-member decoders restore the whole input; each member is what a one-shot call on that batch gives; and the
    reference's own decoder, which reads only the first member (src/unzip.c:361), returns exactly the first batch."""
    import gzip, io
    B = 4 << 20
    n = 40 * (1 << 20) + 999                     # above the small-input threshold: 10 full batches and a short one
    d = gz.corpus("text", n, seed=5)
    src = tmp_path / "in.bin"; src.write_bytes(d.tobytes())
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "4")
    dst = tmp_path / "out.gz"
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6, members=True)
    m = dst.read_bytes()
    assert bi == n and bo == len(m)
    assert gzip.GzipFile(fileobj=io.BytesIO(m)).read() == d.tobytes()                                       # Python gzip
    assert subprocess.run(["gunzip", "-c"], input=m, stdout=subprocess.PIPE, check=True).stdout == d.tobytes()  # system gunzip
    # member by member
    off, k = 0, 0
    while off < len(m):
        part = d[k * B:(k + 1) * B]
        body = orc.engine_model(part, 6).tobytes()
        assert m[off:off + 10] == gz.gzip_header(6)
        assert m[off + 10:off + 10 + len(body)] == body
        assert m[off + 10 + len(body):off + 18 + len(body)] == gz.gzip_trailer(zlib.crc32(part.tobytes()), len(part))
        off += 18 + len(body); k += 1
    assert off == len(m) and k == 11
    if orc.have_ref():
        assert orc.ref_gunzip(m) == d[:B].tobytes()


def test_device_side_verify(gz, orc, ctx):
    """GZB_F_VERIFY: every fragment is decoded again on the device and compared with the input.  It must
    pass on good output of every kind of block and catch a single corrupted byte anywhere."""
    import torch
    V = gz.F_FINAL | gz.F_VERIFY
    for kind, n in (("text", 9 * gz.CHUNK + 77), ("binary", 3 * gz.CHUNK), ("random", 2 * gz.CHUNK + 5), ("mixed", 5 * gz.CHUNK + 1),
                    ("text", 1), ("text", 70000)):
        d = gz.corpus(kind, n, seed=9)
        for level in (1, 6, 9):
            raw, r = ctx.deflate_host(d, level, V)
            assert r.status == gz.OK and raw.tobytes() == orc.engine_model(d, level).tobytes()
        raw, r = ctx.deflate_host(d, 6, V | gz.F_RSYNCABLE)
        assert zlib.decompress(raw.tobytes(), -15) == d.tobytes()
    # shards with a halo, non-final calls
    n = 6 * gz.CHUNK + 123
    d = gz.corpus("text", n, seed=10); t = torch.from_numpy(d).cuda()
    out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    r = ctx.deflate_device(t[2 * gz.CHUNK:], n - 2 * gz.CHUNK, out, 6, flags=V, halo=gz.DICT)
    r = ctx.deflate_device(t, 2 * gz.CHUNK, out, 6, flags=gz.F_VERIFY)
    # a corrupted byte is caught wherever it is (fault injection exists only in the TEST build of the library,
    # csrc/build/libgzb200_testing.so: the shipped one has no such hook)
    import ctypes
    T = ctypes.CDLL(os.path.join(os.path.dirname(gz.LIB_PATH), "csrc", "build", "libgzb200_testing.so"))
    assert not hasattr(gz.lib(), "gzb_test_corrupt_at")
    vp, u64 = ctypes.c_void_p, ctypes.c_uint64
    T.gzb_create.argtypes = [ctypes.c_int, ctypes.POINTER(vp)]; T.gzb_destroy.argtypes = [vp]; T.gzb_destroy.restype = None
    T.gzb_test_corrupt_at.argtypes = [vp, ctypes.c_longlong]
    T.gzb_bound.argtypes = [u64, ctypes.c_uint32]; T.gzb_bound.restype = u64
    T.gzb_deflate_host.argtypes = [vp, ctypes.c_int, ctypes.c_uint32, vp, u64, vp, vp, u64, ctypes.POINTER(gz.Result)]
    h = vp(); assert T.gzb_create(0, ctypes.byref(h)) == 0
    d = gz.corpus("text", 4 * gz.CHUNK, seed=11)
    out = np.empty(T.gzb_bound(d.size, 0), dtype=np.uint8); r = gz.Result()
    def run():
        return T.gzb_deflate_host(h, 6, V, d.ctypes.data, d.size, None, out.ctypes.data, out.size, ctypes.byref(r))
    assert run() == 0
    raw = out[: r.out_bytes].copy()
    assert raw.tobytes() == orc.engine_model(d, 6).tobytes()
    for at in (0, 5, raw.size // 3, raw.size // 2, raw.size - 2, 1 << 40):
        if at < raw.size and raw[at] == 0xA5:
            continue
        T.gzb_test_corrupt_at(h, at)
        assert run() == (gz.E_VERIFY if at < raw.size else 0), at        # an offset beyond the bound is ignored
    T.gzb_test_corrupt_at(h, -1)
    assert run() == 0 and out[: r.out_bytes].tobytes() == raw.tobytes()
    T.gzb_destroy(h)


def test_more_than_4_gib_in_one_call(gz, ctx):
    """64-bit offsets everywhere and ISIZE mod 2^32 (src/parallel.c:60,545): 4 GiB + 3 chunks + 11 bytes in one device call at
    -6, checked over its whole length twice: by the device-side round trip (k_verify) and by zlib, which decodes the complete
    body -- the chunks on either side of the 4 GiB line included -- and must give back every input byte."""
    import torch
    n = (4 << 30) + 3 * gz.CHUNK + 11
    d = gz.corpus("text", n, seed=13)
    t = torch.from_numpy(d).cuda()
    out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    r = ctx.deflate_device(t, n, out, 6, flags=gz.F_FINAL | gz.F_VERIFY)
    assert r.in_bytes == n and r.crc32 == zlib.crc32(memoryview(d))
    assert gz.gzip_trailer(r.crc32, r.in_bytes)[4:] == (n % (1 << 32)).to_bytes(4, "little")
    raw = out[: r.out_bytes].cpu().numpy()
    del t, out
    z, pos, step = zlib.decompressobj(-15), 0, 32 << 20
    for off in range(0, raw.size, step):
        back = z.decompress(memoryview(raw[off:off + step]))
        assert back == d[pos:pos + len(back)].tobytes(), f"decoded bytes differ somewhere in [{pos}, {pos + len(back)})"
        pos += len(back)
    back = z.flush()
    assert back == d[pos:pos + len(back)].tobytes() and pos + len(back) == n and z.eof
