"""Freeze golden vectors from the reference itself (oracle/_ref/refgzip, built from /root/reference by oracle/Makefile).
Run in the build container; the output (tests/golden/golden.json) is committed so the checks travel to the GPU box."""
import hashlib, json, os, sys, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gzippier_b200 as gz
import oracle_lib as orc

out = {"generator": "tools/make_golden.py", "reference": "wfehrnstrom/gzippier src/zip.c + src/parallel.c + lib/zlib 1.2.11 via oracle/_ref/refgzip",
       "seed": 1234, "corpora": {}, "known_answers": {}}
N = 4 << 20
for kind in ("text", "zipf", "binary", "mixed", "random"):
    d = gz.corpus(kind, N, seed=1234)
    e = {"n": N, "sha256": hashlib.sha256(d.tobytes()).hexdigest(), "crc32": zlib.crc32(d.tobytes()), "parallel": {}, "single": {}}
    for level in range(1, 10):
        m = orc.ref_gzip(d, level, threads=4)
        e["parallel"][str(level)] = {"size": len(m), "head": m[:10].hex(), "tail": m[-8:].hex()}
    for level in (1, 6, 9):
        m = orc.ref_gzip(d, level)
        e["single"][str(level)] = {"size": len(m), "head": m[:10].hex(), "tail": m[-8:].hex()}
    e["rsyncable_parallel_6"] = len(orc.ref_gzip(d, 6, threads=4, rsyncable=True))
    out["corpora"][kind] = e
    print(kind, e["parallel"]["6"], flush=True)
# byte-exact known answers (SURVEY.md Appendix A.1 probes, re-generated here)
out["known_answers"]["parallel_nl_j4_L6"] = orc.ref_gzip(b"parallel\n", 6, threads=4).hex()
out["known_answers"]["empty_single_L6"] = orc.ref_gzip(b"", 6).hex()
out["known_answers"]["memcpy_abuse_L6_j2_size"] = len(orc.ref_gzip(b"wxy" + b"%032767d" % 0, 6, threads=2))
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "golden.json"), "w"), indent=1)
