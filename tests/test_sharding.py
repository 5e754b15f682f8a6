"""Host-side multi-GPU logic on CPU: world_size-2 gloo run of bench.py's shard plan (contiguous chunk ranges,
32 KiB halo, crc32_combine, ordered concatenation) -- no data-path collective."""
import os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, zlib
sys.path.insert(0, os.environ["GZB_ROOT"]); sys.path.insert(0, os.path.join(os.environ["GZB_ROOT"], "tests"))
import numpy as np, torch, torch.distributed as dist
import gzippier_b200 as gz, oracle_lib as orc
from gzippier_b200.sharding import shard_plan
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 11 * gz.CHUNK + 999
lo, hi, halo = shard_plan(n, rank, world)
assert lo % gz.CHUNK == 0 and (halo == 0) == (lo == 0)
d = gz.corpus("zipf", n, seed=5)                       # every rank can regenerate any slice
# stand-in for the GPU call in this CPU-only test: the engine's CPU model on the same shard with the same halo
body = orc.engine_model(d[:hi].copy(), 6, halo=halo, final=(rank == world - 1), skip=lo)
crc = zlib.crc32(d[lo:hi].tobytes())
meta = torch.tensor([hi - lo, crc, body.size], dtype=torch.int64)
gathered = [torch.zeros(3, dtype=torch.int64) for _ in range(world)]
dist.all_gather(gathered, meta)                        # measurement/plumbing only; the data path has no collective
sizes = [int(g[2]) for g in gathered]
buf = torch.zeros(max(sizes), dtype=torch.uint8); buf[: body.size] = torch.from_numpy(body.copy())
bufs = [torch.zeros(max(sizes), dtype=torch.uint8) for _ in range(world)]
dist.all_gather(bufs, buf)
if rank == 0:
    whole = b"".join(bufs[r][: sizes[r]].numpy().tobytes() for r in range(world))
    c = 0
    for g in gathered: c = gz.crc32_combine(c, int(g[1]), int(g[0]))
    assert c == zlib.crc32(d.tobytes())
    assert zlib.decompress(whole, -15) == d.tobytes()
    assert whole == orc.engine_model(d, 6).tobytes()  # same bytes as the 1-GPU run: output is independent of the shard count
    print("SHARD_OK")
dist.destroy_process_group()
'''


def test_two_rank_gloo(tmp_path):
    w = tmp_path / "worker.py"; w.write_text(WORKER)
    env = dict(os.environ, GZB_ROOT=ROOT)
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29613", str(w)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300)
    assert p.returncode == 0 and b"SHARD_OK" in p.stdout, p.stdout.decode(errors="replace")[-2000:]


def test_shard_plan_properties():
    sys.path.insert(0, ROOT)
    from gzippier_b200.sharding import shard_plan
    CH = 131072
    for n in (0, 1, CH, CH + 1, 7 * CH, 64 * CH + 5, (1 << 32) + 3):
        for world in (1, 2, 3, 4, 8):
            cover = 0
            for r in range(world):
                lo, hi, halo = shard_plan(n, r, world)
                assert lo == cover and hi >= lo and (lo % CH == 0 or lo == n)
                if hi > lo: assert halo == (32768 if lo else 0)
                cover = hi
            assert cover == n
