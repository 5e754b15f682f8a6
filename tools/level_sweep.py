"""k_lz77 time per level on one corpus (device-resident): python tools/level_sweep.py [MiB] [kind]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 1024 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "text"
ctx = gz.Context(0)
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
for level in range(1, 10):
    ctx.deflate_device(t_in, n, t_out, level)
    r = ctx.deflate_device(t_in, n, t_out, level)
    m = ctx.stage_ms()
    print(f"L{level} {kind}: lz77 {m['lz77']:.2f} ms, total {m['total']:.2f} ms = {n / m['total'] / 1e6:.2f} GB/s, ratio {n / r.out_bytes:.3f}", flush=True)
