"""Small cases for compute-sanitizer (one tool per gpurun call): every k_lz77 instantiation (parser group of 8 / 16 warps,
helpers on / off), stored chunks next to searched ones, a partial last tile, the host path, rsyncable chunks.
usage: compute-sanitizer --tool racecheck|synccheck|memcheck python tools/sanitize_case.py"""
import os, sys, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import gzippier_b200 as gz
cases = []
rnd = gz.corpus("random", 2 * gz.CHUNK, seed=3)
cases.append(("text+random+text", np.concatenate([gz.corpus("text", gz.CHUNK + 777, seed=1)[: gz.CHUNK], rnd, gz.corpus("text", gz.CHUNK + 777, seed=2)])))
cases.append(("mixed, partial tile", gz.corpus("mixed", 3 * gz.CHUNK + 1234, seed=4)))
cases.append(("tiny", gz.corpus("text", 100, seed=5)))
for env in ({}, {"GZB_PW": "16", "GZB_HELP": "3"}, {"GZB_PW": "8", "GZB_HELP": "3", "GZB_ROLEMAP": "0"}, {"GZB_FORCE_RUN": "2"}):
    for k in ("GZB_PW", "GZB_HELP", "GZB_ROLEMAP", "GZB_FORCE_RUN"): os.environ.pop(k, None)
    os.environ.update(env)
    ctx = gz.Context(0)
    for name, d in cases:
        for level, flags in ((1, gz.F_FINAL | gz.F_VERIFY), (6, gz.F_FINAL | gz.F_VERIFY), (9, gz.F_FINAL), (6, gz.F_FINAL | gz.F_RSYNCABLE)):
            raw, r = ctx.deflate_host(d, level, flags)
            assert zlib.decompress(raw.tobytes(), -15) == d.tobytes() and r.crc32 == zlib.crc32(d.tobytes())
    ctx.close()
    print("ok", env, flush=True)
