"""gzb_zip_fd file -> file on tmpfs under several settings of its knobs (one corpus file, written once).
usage: python tools/zipfd_sweep.py [MiB] [n_gpus] ["K=V,K=V;K=V;..."]   (each ;-separated group is one setting)"""
import os, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ngpu = int(sys.argv[2]) if len(sys.argv) > 2 else 1
groups = sys.argv[3].split(";") if len(sys.argv) > 3 else [""]
n = mib << 20
src, dst = "/dev/shm/gzb_in.bin", "/dev/shm/gzb_out.gz"
with open(src, "wb") as f:
    crc = 0
    for off in range(0, n, 1 << 30):
        d = gz.corpus("text", min(1 << 30, n - off), offset=off); f.write(memoryview(d)); crc = zlib.crc32(memoryview(d), crc)
KNOBS = ["GZB_ZIP_WORKERS_PER_GPU", "GZB_ZIP_IO_THREADS", "GZB_ZIP_IO_PIECE_KB", "GZB_ZIP_BATCH_MB", "GZB_ZIP_NO_DIRECT", "GZB_ZIP_VERBOSE", "GZB_ZIP_WRITERS", "GZB_ZIP_NO_MMAP", "GZB_ZIP_FILLERS"]
for g in groups:
    for k in KNOBS: os.environ.pop(k, None)
    for kv in filter(None, g.split(",")):
        k, v = kv.split("="); os.environ["GZB_ZIP_" + k if not k.startswith("GZB_") else k] = v
    best, each = 0.0, []
    for rep in range(int(os.environ.get("ZIPFD_REPS", "3"))):
        with open(src, "rb") as fi, open(dst, "wb") as fo:
            t = time.time(); bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=6, n_gpus=ngpu); dt = time.time() - t
        each.append(round(bi / dt / 1e9, 2))
        if rep: best = max(best, bi / dt / 1e9)
    with open(dst, "rb") as f:
        f.seek(-8, 2); tr = f.read(8)
    ok = int.from_bytes(tr[:4], "little") == crc and int.from_bytes(tr[4:], "little") == n % (1 << 32)
    print(f"[{g or 'defaults'}] {ngpu} GPU, {mib} MiB: {best:6.2f} GB/s warm (best of 2 after the first), {bo} bytes, trailer {'ok' if ok else 'BAD'}; every call: {each}", flush=True)
os.unlink(src); os.unlink(dst)
