"""Stall-reason totals per code region of one kernel from `ncu -i rep --page source --csv` output.
usage: ncu_stalls.py source.csv lo hi   (instruction index range; default all)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 10 ** 9
cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
top = []
n = 0
for r in rows[2:]:
    if len(r) < 5 or r[0] == "Kernel Name":
        break
    if r[0] == "Address":
        continue
    if lo <= n < hi:
        for i in cols:
            try:
                tot[hdr[i]] += int(r[i])
            except ValueError:
                pass
        try:
            top.append((int(r[hdr.index("Warp Stall Sampling (All Samples)")]), n, r[1].strip()[:90],
                        {hdr[i]: int(r[i]) for i in cols if r[i] not in ("", "0")}))
        except ValueError:
            pass
    n += 1
s = sum(tot.values())
for k, v in tot.most_common():
    if v:
        print("%-28s %6d %5.1f%%" % (k, v, 100.0 * v / max(s, 1)))
for t in sorted(top, reverse=True)[:14]:
    print(t)
