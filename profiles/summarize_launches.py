"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python profiles/summarize_launches.py <launches.csv> "<header comment>" """
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr, data = None, []
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr:
        data.append(dict(zip(hdr, r)))
agg = collections.defaultdict(lambda: [0, 0.0])
for d in data:
    n = d["Kernel Name"].split("(")[0][:60]
    v = float(d["Metric Value"].replace(",", ""))
    u = d["Metric Unit"]
    ms = v / 1e6 if u in ("ns", "nsecond") else (v / 1e3 if u.startswith("us") else v)
    agg[n][0] += 1
    agg[n][1] += ms
tot = sum(v[1] for v in agg.values())
if len(sys.argv) > 2:
    print("# " + sys.argv[2])
print("# per-launch times are cold-cache/serialised: compare SHARES with bench.py's stage timing, not absolutes")
for n, (c, ms) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{n:62s} launches={c:4d} total={ms:9.3f} ms share={100 * ms / tot:5.1f}%")
