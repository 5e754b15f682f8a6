"""gzippier_b200 -- Python (ctypes) binding of the C ABI in include/gzb200.h.

Thin plumbing only: device memory and streams come from torch, the engine is
gzippier_b200/libgzb200.so (hand-written sm_100a kernels).  There is no CPU
fallback: if the library is missing or no B200 is present, calls raise.
"""
from __future__ import annotations
import ctypes, os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgzb200.so")
CORPUS_LIB_PATH = os.path.join(_HERE, "libgzbcorpus.so")     # input generator for bench / tests, separate from the engine

OK, E_ARG, E_LEVEL, E_NOMEM, E_CUDA, E_NO_DEVICE, E_OUT_FULL, E_IO, E_VERIFY = 0, -1, -2, -3, -4, -5, -6, -7, -8
E_TOO_LARGE = -9
NOALLOC_MAX_BYTES = 8 << 30            # gzb_crc32_device_noalloc
F_FINAL, F_RSYNCABLE, F_VERIFY = 1, 2, 4
CHUNK, DICT, RSYNC_CHUNK = 131072, 32768, 16384
CORPUS = {"text": 0, "zipf": 1, "binary": 2, "random": 3, "mixed": 4}


class Result(ctypes.Structure):
    _fields_ = [("in_bytes", ctypes.c_uint64), ("out_bytes", ctypes.c_uint64), ("crc32", ctypes.c_uint32),
                ("n_chunks", ctypes.c_uint32), ("n_blocks", ctypes.c_uint32), ("status", ctypes.c_int32)]


class StageMs(ctypes.Structure):
    _fields_ = [(k, ctypes.c_float) for k in ("crc", "lz77", "huffman", "layout", "pack", "total")]


class ZipOpts(ctypes.Structure):
    _fields_ = [("level", ctypes.c_int), ("rsyncable", ctypes.c_int), ("n_gpus", ctypes.c_int), ("name", ctypes.c_char_p),
                ("mtime", ctypes.c_uint32), ("bytes_in", ctypes.c_uint64), ("bytes_out", ctypes.c_uint64),
                ("members", ctypes.c_int)]


class Shard(ctypes.Structure):
    _fields_ = [(k, ctypes.c_uint64) for k in ("in_off", "in_bytes", "out_off", "out_bytes")]


class GzbError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        super().__init__(f"gzb200 error {code}: {detail}")


_lib = None


def lib():
    """Load libgzb200.so (fails loudly if it has not been built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GzbError(E_NO_DEVICE, f"{LIB_PATH} not built -- run `make -C gzippier_b200/csrc`")
        L = ctypes.CDLL(LIB_PATH)
        vp, u64, u32, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int
        L.gzb_create.argtypes = [i32, ctypes.POINTER(vp)]
        L.gzb_destroy.argtypes = [vp]; L.gzb_destroy.restype = None
        L.gzb_strerror.argtypes = [i32]; L.gzb_strerror.restype = ctypes.c_char_p
        L.gzb_last_error.argtypes = [vp]; L.gzb_last_error.restype = ctypes.c_char_p
        L.gzb_device_count.restype = i32
        L.gzb_bound.argtypes = [u64, u32]; L.gzb_bound.restype = u64
        L.gzb_deflate_device.argtypes = [vp, i32, u32, vp, u64, u32, vp, u64, ctypes.POINTER(Result), vp]
        L.gzb_deflate_device_async.argtypes = [vp, i32, u32, vp, u64, u32, vp, u64, vp]
        L.gzb_deflate_wait.argtypes = [vp, ctypes.POINTER(Result)]
        L.gzb_last_stage_ms.argtypes = [vp, ctypes.POINTER(StageMs)]
        L.gzb_deflate_host.argtypes = [vp, i32, u32, vp, u64, vp, vp, u64, ctypes.POINTER(Result)]
        L.gzb_crc32_device.argtypes = [vp, u32, vp, u64, ctypes.POINTER(u32), vp]
        L.gzb_crc32_host.argtypes = [vp, u32, vp, u64, ctypes.POINTER(u32)]
        L.gzb_crc32_device_noalloc.argtypes = [u32, vp, u64, vp, vp]
        L.gzb_crc32_combine.argtypes = [u32, u32, u64]; L.gzb_crc32_combine.restype = u32
        L.gzb_gzip_header.argtypes = [i32, u32, ctypes.c_char_p, vp, ctypes.c_size_t]
        L.gzb_gzip_trailer.argtypes = [u32, u64, vp, ctypes.c_size_t]
        L.gzb_zip_fd.argtypes = [i32, i32, ctypes.POINTER(ZipOpts)]
        L.gzb_zip_release.restype = None
        L.gzb_deflate_host_multi.argtypes = [ctypes.POINTER(vp), i32, i32, u32, vp, u64, vp, vp, u64, ctypes.POINTER(Result), ctypes.POINTER(Shard), i32]
        _lib = L
    return _lib


def crc32_device_noalloc(d_in, n: int, d_crc_out, crc: int = 0, stream: int = 0) -> None:
    """gzb_crc32_device_noalloc (include/gzb200.h): CRC-32 of n device bytes into the device word d_crc_out, on the current
    device, with no allocation.  Enqueues and returns; n <= NOALLOC_MAX_BYTES.  No counterpart in the reference."""
    rc = lib().gzb_crc32_device_noalloc(crc, _ptr(d_in), n, _ptr(d_crc_out), stream or None)
    if rc != 0:
        raise GzbError(rc, lib().gzb_strerror(rc).decode())


def bound(n: int, flags: int = 0) -> int:
    return int(lib().gzb_bound(n, flags))


def crc32_combine(crc1: int, crc2: int, len2: int) -> int:
    return int(lib().gzb_crc32_combine(crc1, crc2, len2))


def gzip_header(level: int, mtime: int = 0, name: bytes | None = None) -> bytes:
    buf = ctypes.create_string_buffer(10 + (len(name) + 1 if name else 0))
    n = lib().gzb_gzip_header(level, mtime, name, buf, len(buf))
    if n < 0:
        raise GzbError(n)
    return buf.raw[:n]


def gzip_trailer(crc: int, isize: int) -> bytes:
    buf = ctypes.create_string_buffer(8)
    lib().gzb_gzip_trailer(crc, isize, buf, 8)
    return buf.raw


_corpus_lib = None


def corpus(kind: str, n: int, seed: int = 1234, offset: int = 0, threads: int = 0, out=None):
    """n bytes of the named synthetic corpus as a numpy uint8 array (host only; libgzbcorpus.so, which is not the engine:
    a process that only calls this maps nothing of libgzb200.so).  `out`: fill this uint8 array instead of a new one."""
    global _corpus_lib
    import numpy as np
    if _corpus_lib is None:
        C = ctypes.CDLL(CORPUS_LIB_PATH)
        C.gzb_corpus_fill.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int]
        _corpus_lib = C
    if out is None:
        out = np.empty(n, dtype=np.uint8)
    assert out.dtype == np.uint8 and out.size >= n and out.flags["C_CONTIGUOUS"]
    rc = _corpus_lib.gzb_corpus_fill(CORPUS[kind], seed, offset, out.ctypes.data, n, threads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError(f"gzb_corpus_fill({kind}) failed")
    return out[:n]


class Context:
    """One engine context on one GPU (gzb_create / gzb_destroy)."""

    def __init__(self, device: int = 0):
        self._h = ctypes.c_void_p()
        rc = lib().gzb_create(device, ctypes.byref(self._h))
        if rc != OK:
            self._h = None
            raise GzbError(rc, lib().gzb_strerror(rc).decode())
        self.device = device

    def close(self):
        if self._h:
            lib().gzb_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc != OK:
            raise GzbError(rc, lib().gzb_last_error(self._h).decode())

    # ---- device-resident ----
    def deflate_device(self, d_in, n: int, d_out, level: int = 6, flags: int = F_FINAL, halo: int = 0, stream: int = 0) -> Result:
        """d_in/d_out: torch uint8 CUDA tensors (or raw device pointers).  Returns Result."""
        r = Result()
        self._check(lib().gzb_deflate_device(self._h, level, flags, _ptr(d_in), n, halo, _ptr(d_out), _cap(d_out), ctypes.byref(r), stream or None))
        return r

    def deflate_device_async(self, d_in, n: int, d_out, level: int = 6, flags: int = F_FINAL, halo: int = 0, stream: int = 0):
        self._check(lib().gzb_deflate_device_async(self._h, level, flags, _ptr(d_in), n, halo, _ptr(d_out), _cap(d_out), stream or None))

    def wait(self) -> Result:
        r = Result()
        self._check(lib().gzb_deflate_wait(self._h, ctypes.byref(r)))
        return r

    def stage_ms(self) -> dict:
        m = StageMs()
        self._check(lib().gzb_last_stage_ms(self._h, ctypes.byref(m)))
        return {k: getattr(m, k) for k, _ in StageMs._fields_}

    # ---- host buffers ----
    def deflate_host(self, data, level: int = 6, flags: int = F_FINAL, dict_bytes=None, out=None):
        """data: bytes-like / numpy uint8.  Returns (raw deflate bytes as numpy view, Result)."""
        import numpy as np
        a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        cap = bound(a.size, flags)
        if out is None:
            out = np.empty(cap, dtype=np.uint8)
        d = None
        if dict_bytes is not None:
            d = np.frombuffer(dict_bytes, dtype=np.uint8) if not isinstance(dict_bytes, np.ndarray) else dict_bytes
            assert d.size == DICT
        r = Result()
        self._check(lib().gzb_deflate_host(self._h, level, flags, a.ctypes.data if a.size else None, a.size,
                                           d.ctypes.data if d is not None else None, out.ctypes.data, out.size, ctypes.byref(r)))
        return out[: r.out_bytes], r

    def crc32_host(self, data, crc: int = 0) -> int:
        import numpy as np
        a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        o = ctypes.c_uint32()
        self._check(lib().gzb_crc32_host(self._h, crc, a.ctypes.data if a.size else None, a.size, ctypes.byref(o)))
        return o.value

    def crc32_device(self, d_in, n: int, crc: int = 0, stream: int = 0) -> int:
        o = ctypes.c_uint32()
        self._check(lib().gzb_crc32_device(self._h, crc, _ptr(d_in), n, ctypes.byref(o), stream or None))
        return o.value

    def gzip(self, data, level: int = 6, rsyncable: bool = False) -> bytes:
        """One complete gzip member for `data` (header + body + trailer), like `gzip -c`."""
        body, r = self.deflate_host(data, level, F_FINAL | (F_RSYNCABLE if rsyncable else 0))
        return gzip_header(level) + body.tobytes() + gzip_trailer(r.crc32, r.in_bytes)


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


def _cap(t):
    if isinstance(t, int) or t is None:
        raise TypeError("output must be a tensor so that its capacity is known")
    return t.numel() * t.element_size()


def zip_fd(ifd: int, ofd: int, level: int = 6, rsyncable: bool = False, n_gpus: int = 1, name: bytes | None = None, mtime: int = 0,
           members: bool = False):
    """Mirror of the reference's parallel_zip(): read ifd to EOF, write one gzip member to ofd
    (members=True: one member per batch, which the reference's own decoder reads only the first of)."""
    o = ZipOpts(level, int(rsyncable), n_gpus, name, mtime, 0, 0, int(members))
    rc = lib().gzb_zip_fd(ifd, ofd, ctypes.byref(o))
    if rc != OK:
        raise GzbError(rc, lib().gzb_strerror(rc).decode())
    return o.bytes_in, o.bytes_out


def zip_release():
    """Free the contexts and page-locked slots gzb_zip_fd keeps between calls."""
    lib().gzb_zip_release()


def deflate_host_multi(contexts, data, level: int = 6, flags: int = F_FINAL, dict_bytes=None, out=None, compact: bool = True):
    """gzb_deflate_host_multi: `data` sharded over the given contexts (one per GPU).  -> (raw bytes view, Result, shards)."""
    import numpy as np
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    cap = bound(a.size, flags) + 64 * len(contexts)
    if out is None:
        out = np.empty(cap, dtype=np.uint8)
    d = None
    if dict_bytes is not None:
        d = np.frombuffer(dict_bytes, dtype=np.uint8) if not isinstance(dict_bytes, np.ndarray) else dict_bytes
    hs = (ctypes.c_void_p * len(contexts))(*[c._h for c in contexts])
    r = Result(); sh = (Shard * len(contexts))()
    rc = lib().gzb_deflate_host_multi(hs, len(contexts), level, flags, a.ctypes.data if a.size else None, a.size,
                                      d.ctypes.data if d is not None else None, out.ctypes.data, out.size, ctypes.byref(r), sh, int(compact))
    if rc != OK:
        raise GzbError(rc, lib().gzb_last_error(contexts[0]._h).decode())
    return (out[: r.out_bytes] if compact else out), r, [(s.in_off, s.in_bytes, s.out_off, s.out_bytes) for s in sh]
