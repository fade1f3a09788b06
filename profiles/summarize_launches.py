"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel name.
usage: python profiles/summarize_launches.py file.csv [skip_substring ...]"""
import collections, csv, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
skip = sys.argv[2:]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if any(s in r[ki] for s in skip):
        continue
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    n = r[ki][:100]
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
for n, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%10.1f us %5d launches %5.1f%%  %s" % (v, c, 100 * v / tot, n))
print("%10.1f us total" % tot)
