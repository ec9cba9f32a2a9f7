"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel launches,
total, mean and share.  usage: python tools/launch_summary.py launches.csv [skip_first_n]"""
import csv, sys, collections, re
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    ms = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v if unit in ("ms", "msecond") else v * 1e3
    rows.append((int(r["ID"]), r["Kernel Name"], ms))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = rows[skip:]
agg = collections.OrderedDict()
for _, name, ms in rows:
    name = re.sub(r"\(.*", "", name)[:90]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += ms
tot = sum(a[1] for a in agg.values())
print(f"{len(rows)} launches, {tot:.3f} ms total")
for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ms:10.3f} ms {100 * ms / tot:5.1f}%  n={n:4d}  avg {ms / n:8.4f} ms  {name}")
