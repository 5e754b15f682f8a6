"""gzb_crc32_device_noalloc (include/gzb200.h): the CRC-32 entry point that allocates nothing -- its per-chunk values live
in a file-scope __device__ array -- against gzb_crc32_device (context-owned cudaMalloc scratch) and zlib, bit for bit.
No counterpart in the reference.  Collected last on purpose (test_zz_*)."""
import ctypes, zlib
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dev(torch, a):
    t = torch.empty(max(a.size, 16), dtype=torch.uint8, device="cuda:0")
    if a.size:
        t[: a.size].copy_(torch.from_numpy(a))
    return t


def test_noalloc_crc32_matches_existing_entry_point_twice(gz, ctx):
    import torch
    torch.cuda.set_device(0)
    out = torch.zeros(1, dtype=torch.int32, device="cuda:0")                # the caller's device word
    # two calls in a row share the static array: the second is SHORTER than the first (stale per-chunk values behind it)
    # and then a ragged one, an empty one and a running-value continuation
    sizes = [5 * gz.CHUNK + 777, 2 * gz.CHUNK + 1, 1, 0, 64 * gz.CHUNK, 300 * gz.CHUNK - 13]
    for rep in range(2):
        for k, n in enumerate(sizes):
            d = gz.corpus("mixed", n, seed=11 + k)
            t = _dev(torch, d)
            gz.crc32_device_noalloc(t, n, out)
            got = out.cpu().numpy().view(np.uint32)[0]
            want = ctx.crc32_device(t, n)
            assert got == want == zlib.crc32(d.tobytes()), (rep, n, hex(got), hex(want))
    # running value: crc(A || B) from crc(A)
    d = gz.corpus("text", 3 * gz.CHUNK + 5, seed=2)
    cut = gz.CHUNK + 16 * 7
    t = _dev(torch, d)
    gz.crc32_device_noalloc(t, cut, out)
    first = int(out.cpu().numpy().view(np.uint32)[0])
    gz.crc32_device_noalloc(t[cut:], d.size - cut, out, crc=first)
    assert out.cpu().numpy().view(np.uint32)[0] == ctx.crc32_device(t[cut:], d.size - cut, crc=first) == zlib.crc32(d.tobytes())


def test_noalloc_allocates_nothing_and_rejects_oversize(gz):
    import torch
    torch.cuda.set_device(0)
    n = 40 * gz.CHUNK + 3
    d = gz.corpus("binary", n, seed=5)
    t = _dev(torch, d)
    out = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    gz.crc32_device_noalloc(t, n, out)                                      # set-up for this device has happened by now
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        gz.crc32_device_noalloc(t, n, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0                                                   # no cudaMalloc behind the call
    assert out.cpu().numpy().view(np.uint32)[0] == zlib.crc32(d.tobytes())
    # beyond the static scratch: a clear error, nothing launched, the output word untouched
    out.fill_(0x1234)
    with pytest.raises(gz.GzbError) as e:
        gz.crc32_device_noalloc(t, gz.NOALLOC_MAX_BYTES + 1, out)
    assert e.value.code == gz.E_TOO_LARGE and "GZB_NOALLOC_MAX_BYTES" in str(e.value)
    torch.cuda.synchronize()
    assert int(out.cpu()[0]) == 0x1234
