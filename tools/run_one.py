"""One device-resident compress of N MiB at a level (used under ncu)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = int(sys.argv[1]) << 20; level = int(sys.argv[2]); kind = sys.argv[3] if len(sys.argv) > 3 else "text"
ctx = gz.Context(0)
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
for _ in range(2):
    r = ctx.deflate_device(t_in, n, t_out, level)
print(n, r.out_bytes, ctx.stage_ms())
