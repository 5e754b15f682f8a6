"""Cycle counters of k_lz77's stations (needs the profiling build: make -C gzippier_b200/csrc LZPROF=1).
python tools/lzprof.py [MiB] [kind] [levels]"""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz, torch
ctx = gz.Context(0); L = gz.lib()
n = int(sys.argv[1]) << 20 if len(sys.argv) > 1 else 64 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "text"
levels = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1, 6, 9]
d = gz.corpus(kind, n); t_in = torch.from_numpy(d).cuda(); t_out = torch.empty(gz.bound(n), dtype=torch.uint8, device="cuda")
# clock64 around the work of one warp per station; the stations run beside one another, so these are latencies under contention
for level in levels:
    ctx.deflate_device(t_in, n, t_out, level)
    buf = (ctypes.c_ulonglong * 16)(); L.gzb_debug_lzprof(ctx._h, buf, 1)
    r = ctx.deflate_device(t_in, n, t_out, level)
    L.gzb_debug_lzprof(ctx._h, buf, 1)
    tiles = buf[7] or 1
    print(f"L{level} {kind} lz77 {ctx.stage_ms()['lz77']:.2f} ms ratio {n/r.out_bytes:.3f} | cycles per tile: "
          f"insert6={buf[8]/tiles:.0f} insert3={buf[9]/tiles:.0f} insert4={buf[10]/tiles:.0f} parse+emit={buf[11]/tiles:.0f}"
          f" | search warp 28: search={buf[0]/tiles:.0f} waits: counter slot={buf[1]/tiles:.0f} tile inserted={buf[2]/tiles:.0f} tile parsed={buf[3]/tiles:.0f}"
          f" | parse phases (thread 0): window exits={buf[12]/tiles:.0f} hop={buf[13]/tiles:.0f} marks={buf[14]/tiles:.0f} prefix+emit={buf[15]/tiles:.0f}", flush=True)
