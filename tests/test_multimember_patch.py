"""SURVEY 8(f4), second half: integration/unzip_multimember.patch makes the reference's decoder read EVERY member of a gzip
file (its inflateGZIP stops after the first, src/unzip.c:361 -- SURVEY F3; the reference's own tests/mixed:43-45 expects
both).  oracle/_ref/refgzip_mm is the reference with the patch applied to a copy of its src/unzip.c (oracle/Makefile);
oracle/_ref/refgzip is the unpatched one.  CPU only, except the last test (the engine's member-per-batch mode)."""
import gzip, io, os, subprocess, zlib
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCHED = os.path.join(ROOT, "oracle", "_ref", "refgzip_mm")
STOCK = os.path.join(ROOT, "oracle", "_ref", "refgzip")
CHUNK = 16384                                  # the decoder's read size (src/unzip.c:41)


@pytest.fixture(scope="module", autouse=True)
def _built():
    if not (os.path.exists(PATCHED) and os.path.exists(STOCK)):
        pytest.skip("oracle/_ref/refgzip_mm is built where /root/reference exists (python __graft_entry__.py)")


def decode(exe, data):
    p = subprocess.run([exe, "-d"], input=data, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=300)
    return p.returncode, p.stdout


def ref_member(data, *args):
    # stdin is a regular file: the reference's single-stream compressor ends its output at the first short read of a pipe
    # (src/zip.c:76-92)
    import tempfile
    with tempfile.TemporaryFile() as f:
        f.write(data); f.flush(); f.seek(0)
        return subprocess.run([STOCK, *args], stdin=f, stdout=subprocess.PIPE, check=True).stdout


def test_two_members_like_tests_mixed():
    two = ref_member(b"xxx\n", "-6") + ref_member(b"yyy\n", "-6")           # (echo xxx | gzip; echo yyy | gzip)
    assert decode(PATCHED, two) == (0, b"xxx\nyyy\n")
    assert decode(STOCK, two) == (0, b"xxx\n")                              # the bug the patch is for
    assert subprocess.run(["gunzip", "-c"], input=two, stdout=subprocess.PIPE, check=True).stdout == b"xxx\nyyy\n"


def test_single_member_is_unchanged():
    rng = np.random.default_rng(3)
    for n in (0, 1, 9, CHUNK - 1, CHUNK, CHUNK + 1, 300000):
        data = bytes(rng.integers(97, 101, n, dtype=np.uint8))
        for args in (["-6"], ["-1", "-j", "4"]):
            if n == 0 and "-j" in args:
                continue                                                    # the reference's -j path deadlocks on empty input
            m = ref_member(data, *args)
            assert decode(PATCHED, m) == decode(STOCK, m) == (0, data)


def test_member_boundaries_at_every_offset_of_the_read_buffer():
    """The next member's magic may start anywhere in the 16 KiB read buffer, also on its last byte (one byte of the magic in
    each of two reads) and exactly at its end (nothing left when the first member ends)."""
    rng = np.random.default_rng(5)
    tail = bytes(rng.integers(0, 256, 5000, dtype=np.uint8))
    second = ref_member(tail, "-6")
    for want in (CHUNK - 2, CHUNK - 1, CHUNK, CHUNK + 1, 2 * CHUNK - 1, 2 * CHUNK, 3 * CHUNK + 7):
        # a first member of exactly `want` bytes: stored blocks of random data, padded through the length of its FNAME-less payload
        n = want - 18 - 5 * ((want - 18 + 65534) // 65535)
        while True:
            body = bytes(rng.integers(0, 256, max(n, 0), dtype=np.uint8))
            co = zlib.compressobj(0, zlib.DEFLATED, 31)
            m = co.compress(body) + co.flush()
            if len(m) == want:
                break
            n += want - len(m)
        assert decode(PATCHED, m + second) == (0, body + tail), want
        assert decode(PATCHED, m + second + m) == (0, body + tail + body), want


def test_many_members_and_trailing_zeros():
    rng = np.random.default_rng(7)
    parts = [bytes(rng.integers(65, 70, int(k), dtype=np.uint8)) for k in rng.integers(0, 70000, 40)]
    members = b"".join(ref_member(p, "-6") if p else gzip.compress(p) for p in parts)
    assert decode(PATCHED, members) == (0, b"".join(parts))
    assert gzip.GzipFile(fileobj=io.BytesIO(members)).read() == b"".join(parts)
    assert decode(PATCHED, members + b"\0" * 1000) == (0, b"".join(parts))   # zero padding after the last member is ignored, as before
    rc, _ = decode(PATCHED, members[:-5])                                    # a cut last member is still an error
    assert rc != 0


@pytest.mark.gpu
def test_engine_member_per_batch_mode_decodes_whole_with_the_patched_reference(gz, tmp_path, monkeypatch):
    """gzb_zip_opts.members (first half): one member per batch.  The unpatched reference restores only the first
    batch; with the patch it restores the whole input."""
    B = 4 << 20
    big = gz.corpus("text", 36 * (1 << 20) + 5, seed=9)                      # above gzb_zip_fd's small-input threshold
    src, dst = tmp_path / "in.bin", tmp_path / "out.gz"
    src.write_bytes(big.tobytes())
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "4")
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6, members=True)
    m = dst.read_bytes()
    assert bi == big.size and bo == len(m)
    assert decode(PATCHED, m) == (0, big.tobytes())
    assert decode(STOCK, m) == (0, big[:B].tobytes())
