"""Throughput of the reference-facing host driver gzb_zip_fd (file -> file on tmpfs, and through pipes with the CLI).
usage: python tools/zipfd_bench.py [MiB] [level] [n_gpus]"""
import os, subprocess, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gzippier_b200 as gz
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
level = int(sys.argv[2]) if len(sys.argv) > 2 else 6
ngpu = int(sys.argv[3]) if len(sys.argv) > 3 else 1
n = mib << 20
src, dst = "/dev/shm/gzb_in.bin", "/dev/shm/gzb_out.gz"
d = gz.corpus("text", n)
with open(src, "wb") as f: f.write(memoryview(d))
crc = zlib.crc32(memoryview(d))
for rep in range(3):
    with open(src, "rb") as fi, open(dst, "wb") as fo:
        t = time.time(); bi, bo = gz.zip_fd(fi.fileno(), fo.fileno(), level=level, n_gpus=ngpu); dt = time.time() - t
    print(f"zip_fd file->file  L{level} {ngpu} GPU: {bi / dt / 1e9:6.2f} GB/s  ({bi} -> {bo} bytes, {dt * 1e3:.0f} ms incl. context creation)", flush=True)
m = open(dst, "rb").read()
assert int.from_bytes(m[-8:-4], "little") == crc and int.from_bytes(m[-4:], "little") == n % (1 << 32)
if mib <= 4096:
    t = time.time(); ok = zlib.decompress(m, 31) == d.tobytes(); print("round trip (zlib):", ok, f"{time.time() - t:.1f}s", flush=True)
exe = os.path.join(os.path.dirname(gz.LIB_PATH), "gzb200")
t = time.time(); subprocess.run(f"cat {src} | {exe} -{level} > /dev/null", shell=True, check=True); dt = time.time() - t
print(f"cat | gzb200 -{level} > /dev/null: {n / dt / 1e9:6.2f} GB/s", flush=True)
os.unlink(src); os.unlink(dst)
