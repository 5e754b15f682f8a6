"""Exploration harness for oracle/engine_model.c: ratio of the engine's algorithm vs the
reference (refgzip -j) per level and corpus.  CPU only; not part of the product or the tests."""
import ctypes, subprocess, sys, time, zlib, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
subprocess.check_call(f"gcc -O2 -shared -fPIC -I{ROOT}/include -o /tmp/libcorpus.so {ROOT}/gzippier_b200/csrc/corpus.c -lpthread -lm && "
                      f"gcc -O3 -march=native -shared -fPIC -o /tmp/libmodel.so {ROOT}/oracle/engine_model.c -lpthread", shell=True)
C = ctypes.CDLL('/tmp/libcorpus.so'); M = ctypes.CDLL('/tmp/libmodel.so')
C.gzb_corpus_fill.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int]
class Params(ctypes.Structure):
    _fields_ = [(k, ctypes.c_int) for k in "level good max_lazy nice ref_chain depth lazy_parse maxd hbits block_span chunk dict use_h3 use_h4 scan good_cut".split()]
class Stats(ctypes.Structure):
    _fields_ = [("tokens", ctypes.c_uint64), ("matches", ctypes.c_uint64), ("blocks", ctypes.c_uint64 * 3)]
M.gzm_deflate.restype = ctypes.c_int64
M.gzm_deflate.argtypes = [ctypes.POINTER(Params), ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(Stats)]
def corpus(kind, n, seed=1234):
    b = np.empty(n, dtype=np.uint8); C.gzb_corpus_fill(kind, seed, 0, b.ctypes.data, n, 8); return b
def ref_size(data, level):
    p = subprocess.run([f"{ROOT}/oracle/_ref/refgzip", f"-{level}", "-j", "8"], input=data.tobytes(), stdout=subprocess.PIPE, check=True)
    return len(p.stdout) - 18
def model(data, level, **over):
    P = Params(); M.gzm_level_params(level, ctypes.byref(P))
    for k, v in over.items(): setattr(P, k, v)
    out = np.empty(len(data) + len(data) // 8 + 65536, dtype=np.uint8); st = Stats()
    t = time.time()
    n = M.gzm_deflate(ctypes.byref(P), data.ctypes.data, len(data), 0, 1, out.ctypes.data, len(out), 8, ctypes.byref(st))
    dt = time.time() - t
    assert n > 0
    raw = out[:n].tobytes()
    assert zlib.decompress(raw, -15) == data.tobytes(), "ROUND TRIP FAILED"
    return n, st, dt
if __name__ == "__main__":
    n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 16 << 20
    names = ['text', 'zipf', 'binary', 'random', 'mixed']
    for kind in (0, 1, 2, 4):
        d = corpus(kind, n)
        for level in (1, 6, 9):
            r = ref_size(d, level)
            m, st, dt = model(d, level)
            print(f"{names[kind]:7s} L{level} ref {r:10d} model {m:10d} ratio {m/r:.4f} tok {st.tokens} match {st.matches} blocks {list(st.blocks)} {dt:.1f}s", flush=True)
