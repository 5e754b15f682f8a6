"""k_lz77 scheduling knobs (role layout, helpers, parser group, run length) on one corpus, device-resident; the bytes must not
depend on them.  The knobs are read by gzb_create, so every variant gets its own context.
python tools/variant_sweep.py [MiB] [kind] [levels, e.g. 5,6,9] [variants, e.g. "GZB_ROLEMAP=0;GZB_ROLEMAP=2,GZB_HELP=3"]"""
import hashlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 1024 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "text"
levels = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [6]
KNOBS = ("GZB_ROLEMAP", "GZB_HELP", "GZB_MAX_RUN", "GZB_PW", "GZB_FORCE_RUN")
variants = [dict()]
if len(sys.argv) > 4:
    variants += [dict(kv.split("=") for kv in v.split(",") if kv) for v in sys.argv[4].split(";")]
else:
    variants += [{"GZB_ROLEMAP": r, "GZB_HELP": h} for r in (0, 1, 2) for h in (0, 1, 2, 3)]
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
for level in levels:
    ref = None
    for v in variants:
        for k in KNOBS: os.environ.pop(k, None)
        for k, x in v.items(): os.environ[k] = str(x)
        ctx = gz.Context(0)
        best = 1e9
        for _ in range(3):
            r = ctx.deflate_device(t_in, n, t_out, level)
            m = ctx.stage_ms(); best = min(best, m["lz77"])
        h = hashlib.sha1(t_out[: r.out_bytes].cpu().numpy().tobytes()).hexdigest()[:12]
        ctx.close()
        if ref is None: ref = h
        print(f"L{level} {kind} {v}: lz77 {best:.2f} ms = {n / best / 1e6:.2f} GB/s (lz77 alone), total {m['total']:.2f} ms, sha {h} {'same' if h == ref else 'DIFFERENT'}", flush=True)
