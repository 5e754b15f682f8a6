"""The fd driver (gzippier_b200/host/gzb_zip.c: what stands where the reference's parallel_zip() stood, src/parallel.c:699-834)
on the CPU: built against tests/mock_engine.c, a stand-in that turns every batch into stored deflate blocks, so that the
driver's own work is checked without a GPU -- batching, the 32 KiB handed over as each batch's dictionary (:767-774), order,
the offsets of the file-to-file path, header / trailer / folded CRC (:496-507, :544), the member-per-batch mode, errors.
The mock is test infrastructure (it lives here, the product library never links it; the product has no CPU path)."""
import ctypes, gzip, io, os, subprocess, threading, zlib
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "build")
LIB = os.path.join(BUILD, "libzipmock.so")
MB = 1 << 20


@pytest.fixture(scope="module")
def drv():
    os.makedirs(BUILD, exist_ok=True)
    srcs = [os.path.join(ROOT, "gzippier_b200", "host", "gzb_zip.c"), os.path.join(ROOT, "tests", "mock_engine.c"),
            os.path.join(ROOT, "oracle", "crc32_oracle.c")]
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", LIB, *srcs, "-lpthread"])
    from gzippier_b200 import ZipOpts
    L = ctypes.CDLL(LIB)
    L.gzb_zip_fd.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ZipOpts)]
    L.gzb_zip_release.restype = None
    L.mock_dict_xor.restype = ctypes.c_uint32
    L.mock_dict_count.restype = ctypes.c_long

    def zip_fd(ifd, ofd, level=6, rsyncable=False, n_gpus=1, name=None, mtime=0, members=False):
        o = ZipOpts(level, int(rsyncable), n_gpus, name, mtime, 0, 0, int(members))
        rc = L.gzb_zip_fd(ifd, ofd, ctypes.byref(o))
        return rc, o.bytes_in, o.bytes_out
    L.zip_fd = zip_fd
    yield L
    L.gzb_zip_release()


def data_of(n, seed=1):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, n, dtype=np.uint8).tobytes()


def dict_xor(data, batch, n_batches_with_dict=None):
    x, k = 0, 0
    for off in range(batch, len(data), batch):
        x ^= zlib.crc32(data[off - 32768:off]); k += 1
    return x, k


@pytest.mark.parametrize("env", [{}, {"GZB_ZIP_NO_MMAP": "1"}, {"GZB_ZIP_NO_DIRECT": "1"}, {"GZB_ZIP_WORKERS_PER_GPU": "3", "GZB_ZIP_IO_THREADS": "2"},
                                 {"MOCK_GPUS": "4"}, {"MOCK_GPUS": "3", "GZB_ZIP_NO_DIRECT": "1"}])
def test_file_to_file(drv, tmp_path, monkeypatch, env):
    """Above the small-input threshold (32 MiB): many batches, several workers; the member decodes to the input, the trailer
    is the input's CRC-32 and length, every batch got the 32 KiB before it as its dictionary, the descriptors end where
    sequential I/O would leave them."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "3")
    n = 37 * MB + 4321
    data = data_of(n)
    src, dst = tmp_path / "in", tmp_path / "out.gz"
    src.write_bytes(b"HEAD" * 25 + data)
    drv.gzb_zip_release(); drv.mock_reset()
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        fi.seek(100); fo.write(b"keep"); fo.flush()
        rc, bi, bo = drv.zip_fd(fi.fileno(), fo.fileno(), level=6, n_gpus=int(env.get("MOCK_GPUS", "1")), name=b"in", mtime=77)
        assert rc == 0 and bi == n
        assert os.lseek(fi.fileno(), 0, os.SEEK_CUR) == 100 + n and os.lseek(fo.fileno(), 0, os.SEEK_CUR) == 4 + bo
    m = dst.read_bytes()
    assert m[:4] == b"keep" and len(m) == 4 + bo
    m = m[4:]
    assert m[:10] == bytes([31, 139, 8, 8, 77, 0, 0, 0, 0, 3]) and m[10:13] == b"in\0"
    assert m[-8:] == zlib.crc32(data).to_bytes(4, "little") + (n % (1 << 32)).to_bytes(4, "little")
    assert zlib.decompress(m, 31) == data
    want_xor, want_n = dict_xor(data, 3 * MB)
    assert (drv.mock_dict_xor(), drv.mock_dict_count()) == (want_xor, want_n)


def test_pipes_small_and_empty_inputs(drv, tmp_path, monkeypatch):
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "2")
    for n in (0, 1, 131072, 5 * MB + 7, 33 * MB + 1):
        data = data_of(n, seed=n + 3)
        r_in, w_in = os.pipe(); r_out, w_out = os.pipe()
        got = bytearray()
        t1 = threading.Thread(target=lambda: (os.write(w_in, data) if n < 60000 else os.fdopen(w_in, "wb", closefd=False).write(data), os.close(w_in)))
        t2 = threading.Thread(target=lambda: got.extend(os.fdopen(r_out, "rb").read()))
        t1.start(); t2.start()
        rc, bi, bo = drv.zip_fd(r_in, w_out, level=1)
        os.close(w_out); os.close(r_in); t1.join(); t2.join()
        assert rc == 0 and bi == n and bo == len(got)
        assert got[8] == 4 and zlib.decompress(bytes(got), 31) == data                    # XFL = 4 at -1 (src/parallel.c:43)
        if n == 0:
            assert len(got) == 10 + 5 + 8                                                 # the mock's empty final stored block


def test_member_per_batch_and_rsyncable(drv, tmp_path, monkeypatch):
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "4")
    n = 35 * MB + 99
    data = data_of(n, seed=9)
    src, dst = tmp_path / "in", tmp_path / "out.gz"
    src.write_bytes(data)
    drv.mock_reset()
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        rc, bi, bo = drv.zip_fd(fi.fileno(), fo.fileno(), members=True)
    m = dst.read_bytes()
    assert rc == 0 and bo == len(m) and drv.mock_dict_count() == 0                        # members share no dictionary
    assert gzip.GzipFile(fileobj=io.BytesIO(m)).read() == data
    first = zlib.decompressobj(31); assert first.decompress(m) == data[:4 * MB] and first.unused_data[:3] == b"\x1f\x8b\x08"
    drv.mock_reset()
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        rc, bi, bo = drv.zip_fd(fi.fileno(), fo.fileno(), rsyncable=True)
    assert rc == 0 and drv.mock_dict_count() == 0 and zlib.decompress(dst.read_bytes(), 31) == data   # --rsyncable: no dictionary


@pytest.mark.parametrize("direct", [True, False])
def test_errors_are_reported(drv, tmp_path, monkeypatch, direct):
    """the reference ignores write errors; here an unwritable output, a failing engine call in the middle of the
    stream and a bad level come back as error codes (and nothing hangs)."""
    if not direct:
        monkeypatch.setenv("GZB_ZIP_NO_DIRECT", "1")
    monkeypatch.setenv("GZB_ZIP_BATCH_MB", "2")
    src = tmp_path / "in"; src.write_bytes(data_of(40 * MB, seed=4))
    (tmp_path / "ro").write_bytes(b"")
    with open(src, "rb") as fi, open(tmp_path / "ro", "rb") as fo:
        assert drv.zip_fd(fi.fileno(), fo.fileno())[0] == -7                              # GZB_E_IO
    drv.gzb_zip_release()
    monkeypatch.setenv("MOCK_FAIL_AFTER", "5")
    with open(src, "rb") as fi, open(tmp_path / "o", "wb") as fo:
        assert drv.zip_fd(fi.fileno(), fo.fileno())[0] == -4                              # the engine's code is passed on
    monkeypatch.delenv("MOCK_FAIL_AFTER")
    drv.gzb_zip_release()
    with open(src, "rb") as fi, open(tmp_path / "o", "wb") as fo:
        assert drv.zip_fd(fi.fileno(), fo.fileno(), level=0)[0] == -2                     # GZB_E_LEVEL
        assert drv.zip_fd(fi.fileno(), fo.fileno(), level=6)[0] == 0                      # and the driver works again afterwards
    assert zlib.decompress((tmp_path / "o").read_bytes(), 31) == src.read_bytes()
