"""Cost of GZB_F_VERIFY: device-resident compress with and without the device-side round-trip check."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 4096 << 20
ctx = gz.Context(0)
for kind, level in (("text", 6), ("text", 1), ("random", 1), ("mixed", 9)):
    d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
    res = {}
    for name, fl in (("plain", gz.F_FINAL), ("verify", gz.F_FINAL | gz.F_VERIFY)):
        ctx.deflate_device(t_in, n, t_out, level, flags=fl)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = ctx.deflate_device(t_in, n, t_out, level, flags=fl); e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1)
    dv = res["verify"] - res["plain"]
    print(f"{kind} L{level}: compress {res['plain']:.1f} ms, with verify {res['verify']:.1f} ms -> verify {dv:.1f} ms = {n / dv / 1e6:.1f} GB/s of input", flush=True)
    del t_in, t_out
