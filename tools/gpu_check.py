"""First-contact GPU check: round trip + CRC + byte-equality with the CPU model, then stage timings."""
import ctypes, os, subprocess, sys, time, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gzippier_b200 as gz
subprocess.check_call(f"gcc -O2 -shared -fPIC -o /tmp/libmodel.so {ROOT}/oracle/engine_model.c -lpthread", shell=True)
M = ctypes.CDLL('/tmp/libmodel.so')
class Params(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int) for k in "level good max_lazy nice ref_chain depth lazy_parse maxd hbits block_span chunk dict use_h3 use_h4 scan good_cut".split()]
M.gzm_deflate.restype = ctypes.c_int64
M.gzm_deflate.argtypes = [ctypes.POINTER(Params), ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_void_p]
def model(data, level):
    P = Params(); M.gzm_level_params(level, ctypes.byref(P))
    out = np.empty(len(data) + len(data) // 8 + 65536, dtype=np.uint8)
    n = M.gzm_deflate(ctypes.byref(P), data.ctypes.data, len(data), 0, 1, out.ctypes.data, len(out), 8, None)
    return out[:n]
ctx = gz.Context(0)
bad = 0
cases = [("text", 0), ("text", 1), ("text", 5), ("text", 1000), ("text", 2047), ("text", 2048), ("text", 2049), ("text", 65536), ("text", 131071), ("text", 131072), ("text", 131073),
         ("text", 300000), ("zipf", 1 << 20), ("binary", 1 << 20), ("random", 300000), ("mixed", 4 << 20), ("text", 8 << 20)]
for kind, n in cases:
    d = gz.corpus(kind, n)
    for level in (1, 6, 9) if n <= (1 << 20) else (6,):
        try:
            body, r = ctx.deflate_host(d, level)
        except Exception as e:
            print("FAIL", kind, n, level, e); bad += 1; continue
        raw = body.tobytes()
        ok_crc = r.crc32 == zlib.crc32(d.tobytes())
        try:
            back = zlib.decompress(raw, -15); ok_rt = back == d.tobytes()
        except Exception as e:
            ok_rt = False; back = str(e)
        m = model(d, level).tobytes()
        same = m == raw
        first = next((i for i, (a, b) in enumerate(zip(m, raw)) if a != b), min(len(m), len(raw))) if not same else -1
        print(f"{kind:7s} n={n:9d} L{level} out={len(raw):9d} model={len(m):9d} crc={'ok' if ok_crc else 'BAD'} roundtrip={'ok' if ok_rt else 'BAD'} model_eq={'ok' if same else 'DIFF@%d' % first}", flush=True)
        if not (ok_crc and ok_rt and same):
            bad += 1
            if not ok_rt and isinstance(back, str): print("   inflate:", back)
print("bad:", bad)
if bad == 0 and len(sys.argv) > 1:
    import torch
    n = int(sys.argv[1]) << 20
    d = gz.corpus("text", n)
    t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    for level in (1, 6, 9):
        for it in range(3):
            torch.cuda.synchronize(); t = time.time()
            r = ctx.deflate_device(t_in, n, t_out, level)
            torch.cuda.synchronize(); dt = time.time() - t
        print(f"L{level}: {n/dt/1e9:.2f} GB/s in, ratio {n/r.out_bytes:.3f}, stages {ctx.stage_ms()}", flush=True)
        raw = t_out[: r.out_bytes].cpu().numpy().tobytes()
        assert zlib.decompress(raw, -15) == d.tobytes()
sys.exit(1 if bad else 0)
