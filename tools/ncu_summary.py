"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total time and share.
usage: ncu_summary.py launches.csv [last_n_launches]"""
import collections, csv, sys
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
r = csv.reader(lines)
hdr = next(r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
for row in r:
    v = float(row[vi].replace(",", ""))
    v = v / 1000 if row[ui] == "ns" else (v * 1000 if row[ui] == "ms" else v)
    rows.append((row[ki], v))
if len(sys.argv) > 2:
    rows = rows[-int(sys.argv[2]):]
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in rows:
    k = k.split("(")[0][-70:]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print("%d launches, %.1f us total (per-launch times under ncu are cold-cache and serialised)" % (len(rows), tot))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-72s n=%4d total=%10.1f us share=%5.1f%%" % (k, v[0], v[1], 100 * v[1] / tot))
