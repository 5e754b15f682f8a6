"""Determinism / race stress: every level, several repetitions on the same input, outputs must be identical and pass the
device-side round trip.  python tools/stress.py [MiB] [kind] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
n = (int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 1024 << 20) + 4321
kind = sys.argv[2] if len(sys.argv) > 2 else "mixed"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
ctx = gz.Context(0)
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda()
outs = [torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda") for _ in range(2)]
bad = 0
for level in range(1, 10):
    r0 = ctx.deflate_device(t_in, n, outs[0], level, flags=gz.F_FINAL | gz.F_VERIFY)
    for k in range(reps):
        r = ctx.deflate_device(t_in, n, outs[1], level, flags=gz.F_FINAL | gz.F_VERIFY)
        same = r.out_bytes == r0.out_bytes and r.crc32 == r0.crc32 and bool(torch.equal(outs[0][: r0.out_bytes], outs[1][: r.out_bytes]))
        if not same: bad += 1
    print(f"L{level} {kind}: {n} -> {r0.out_bytes}, {reps} repetitions identical: {bad == 0}", flush=True)
print("bad:", bad)
