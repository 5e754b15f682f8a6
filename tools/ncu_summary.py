"""Summarise one `ncu --set full --import-source on` capture: key metrics + per-source-line instruction shares.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "lts__t_bytes.sum", "sm__cycles_elapsed.max"]
for v in rows[2:]:
    print("== launch ==")
    for i, n in enumerate(h):
        if n in KEYS or n.startswith("smsp__average_warps_issue_stalled") and n.endswith("per_issue_active.ratio"):
            print(f"  {n} [{u[i]}] = {v[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, agg = None, {}
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) > 10 and r[2] == "-":
        try: line, samples, inst = int(r[0]), int(r[6]), int(r[7])
        except ValueError: continue
        a = agg.setdefault((cur, line), [0, 0, r[1].strip()[:100]]); a[0] += inst; a[1] += samples
tot = sum(v[0] for v in agg.values()) or 1; ts = sum(v[1] for v in agg.values()) or 1
print(f"== source lines by warp instructions executed (total {tot}, {ts} pc samples) ==")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
    print(f"  {k[0][:16]:16s}:{k[1]:4d} inst {100*v[0]/tot:5.1f}% samples {100*v[1]/ts:5.1f}%  {v[2]}")
