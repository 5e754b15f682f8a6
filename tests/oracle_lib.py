"""Access to the checker: oracle/liboracle.so (C restatement), oracle/_ref (the reference itself).
Test infrastructure only -- nothing under gzippier_b200/ imports this."""
import ctypes, os, subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
REFGZIP = os.path.join(REF_DIR, "refgzip")
REFZ_SO = os.path.join(REF_DIR, "librefz.so")

_o = None


def lib():
    global _o
    if _o is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
        L = ctypes.CDLL(ORACLE_SO)
        vp, u64, u32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32
        L.orc_crc32.argtypes = [u32, vp, ctypes.c_size_t]; L.orc_crc32.restype = u32
        L.orc_updcrc.argtypes = [u32, vp, ctypes.c_size_t]; L.orc_updcrc.restype = u32
        L.orc_crc32_combine.argtypes = [u32, u32, u64]; L.orc_crc32_combine.restype = u32
        L.orc_gzip_header.argtypes = [ctypes.c_int, u32, ctypes.c_char_p, vp]; L.orc_gzip_header.restype = ctypes.c_size_t
        L.orc_gzip_header_single.argtypes = [ctypes.c_int, vp]; L.orc_gzip_header_single.restype = ctypes.c_size_t
        L.orc_gzip_trailer.argtypes = [u32, u64, vp]; L.orc_gzip_trailer.restype = ctypes.c_size_t
        L.orc_inflate.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]; L.orc_inflate.restype = ctypes.c_int64
        L.gzm_deflate.restype = ctypes.c_int64
        L.gzm_deflate.argtypes = [vp, vp, u64, ctypes.c_int, ctypes.c_int, vp, u64, ctypes.c_int, vp]
        L.gzm_level_params.argtypes = [ctypes.c_int, vp]
        _o = L
    return _o


def _arr(data):
    return data if isinstance(data, np.ndarray) else np.frombuffer(bytes(data), dtype=np.uint8)


def crc32(data, crc=0):
    a = _arr(data)
    return lib().orc_crc32(crc, a.ctypes.data if a.size else None, a.size) if a.size else crc


def crc32_combine(c1, c2, len2):
    return lib().orc_crc32_combine(c1, c2, len2)


def gzip_header(level, mtime=0, name=None):
    buf = ctypes.create_string_buffer(16 + (len(name) if name else 0))
    n = lib().orc_gzip_header(level, mtime, name, buf)
    return buf.raw[:n]


def gzip_header_single(level):
    buf = ctypes.create_string_buffer(16)
    n = lib().orc_gzip_header_single(level, buf)
    return buf.raw[:n]


def gzip_trailer(crc, ulen):
    buf = ctypes.create_string_buffer(8)
    lib().orc_gzip_trailer(crc, ulen, buf)
    return buf.raw


class InflateStats(ctypes.Structure):
    _fields_ = [("blocks", ctypes.c_uint64 * 3), ("literals", ctypes.c_uint64), ("matches", ctypes.c_uint64),
                ("out_bytes", ctypes.c_uint64), ("in_bits", ctypes.c_uint64), ("bfinal_seen", ctypes.c_int)]


def inflate(raw, out_cap, stats=None):
    """Raw-deflate decode with the oracle's own decoder; returns numpy bytes or None on error."""
    a = _arr(raw)
    out = np.empty(max(out_cap, 1), dtype=np.uint8)
    st = InflateStats()
    n = lib().orc_inflate(a.ctypes.data, a.size, out.ctypes.data, out_cap, ctypes.byref(st))
    if stats is not None:
        stats.append(st)
    if n < 0:
        return None
    return out[:n]


class ModelParams(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int) for k in "level good max_lazy nice ref_chain depth lazy_parse maxd hbits block_span chunk dict use_h3 use_h4 scan good_cut".split()]


def engine_model(data, level, halo=0, final=True, rsyncable=False, threads=8, skip=0):
    """Bytes the ENGINE must produce for data[skip:] (data[skip-halo:skip] being the dictionary), from the CPU model."""
    a = _arr(data)
    P = ModelParams(); lib().gzm_level_params(level, ctypes.byref(P))
    if rsyncable:
        P.chunk = 16384; P.dict = 0
    n = a.size - skip
    out = np.empty(n + n // 8 + 65536, dtype=np.uint8)
    r = lib().gzm_deflate(ctypes.byref(P), a.ctypes.data + skip, n, halo, 1 if final else 0, out.ctypes.data, out.size, threads, None)
    assert r >= 0
    return out[:r]


# ---- the reference itself (oracle/_ref) ----
def have_ref():
    return os.path.exists(REFGZIP)


def ref_gzip(data, level=6, threads=0, rsyncable=False):
    cmd = [REFGZIP, f"-{level}"] + (["-j", str(threads)] if threads else []) + (["--rsyncable"] if rsyncable else [])
    # from a regular file, not a pipe: the reference's single-stream loop takes any short read for the end of the input
    # (src/zip.c:76-92: flush = avail_in != CHUNK ? Z_FINISH : ...), so a pipe whose writer is a moment late truncates it
    import tempfile
    with tempfile.TemporaryFile() as f:
        f.write(_arr(data).tobytes()); f.flush(); f.seek(0)
        return subprocess.run(cmd, stdin=f, stdout=subprocess.PIPE, check=True, timeout=600).stdout


def ref_gunzip(member):
    p = subprocess.run([REFGZIP, "-d"], input=bytes(member), stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    if p.returncode != 0:
        raise RuntimeError("reference gzip -d failed: " + p.stderr.decode(errors="replace"))
    return p.stdout


_z = None


def refz():
    """The reference's vendored zlib 1.2.11 as a shared object (crc32, crc32_combine, inflate...)."""
    global _z
    if _z is None:
        Z = ctypes.CDLL(REFZ_SO)
        Z.crc32.argtypes = [ctypes.c_ulong, ctypes.c_void_p, ctypes.c_uint]; Z.crc32.restype = ctypes.c_ulong
        Z.crc32_combine.argtypes = [ctypes.c_ulong, ctypes.c_ulong, ctypes.c_long]; Z.crc32_combine.restype = ctypes.c_ulong
        _z = Z
    return _z
