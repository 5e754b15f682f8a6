"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total, share."""
import csv, collections, sys
lines = open(sys.argv[1]).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    k = r['Kernel Name'].split('(')[0].split('::')[-1]
    agg[k][0] += 1; agg[k][1] += float(r['Metric Value']) / 1e6
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot:.3f} ms total (ncu per-launch times: cold cache, serialised)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:44s} n={v[0]:4d} total {v[1]:10.3f} ms  avg {v[1]/v[0]:9.4f} ms  share {100*v[1]/tot:5.1f}%")
