"""k_lz77 scheduling knobs (role map, spin interval, flow) on one corpus, device-resident; the bytes must not depend on them.
python tools/variant_sweep.py [MiB] [kind] [levels, e.g. 5,6,9]"""
import hashlib, itertools, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 1024 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "text"
levels = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [6]
ctx = gz.Context(0)
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
variants = [dict()]
for level in levels:
    ref = None
    for v in variants:
        for k in ("GZB_ROLEMAP", "GZB_SPIN", "GZB_FLOW", "GZB_MAX_RUN", "GZB_PW"): os.environ.pop(k, None)
        for k, x in v.items(): os.environ[k] = str(x)
        best = 1e9
        for _ in range(3):
            r = ctx.deflate_device(t_in, n, t_out, level)
            m = ctx.stage_ms(); best = min(best, m["lz77"])
        h = hashlib.sha1(t_out[: r.out_bytes].cpu().numpy().tobytes()).hexdigest()[:12]
        if ref is None: ref = h
        print(f"L{level} {kind} {v}: lz77 {best:.2f} ms = {n / best / 1e6:.2f} GB/s (lz77 alone), total {m['total']:.2f} ms, sha {h} {'same' if h == ref else 'DIFFERENT'}", flush=True)
