"""per-kernel totals of an ncu launch list (gpu__time_duration.sum csv)"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) < 15:
        continue
    name = r[4].split("(")[0]
    agg[name][0] += 1
    agg[name][1] += float(r[-1])
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 24]:
    print("%-44s n=%4d total %9.1f us (%4.1f%%) mean %8.2f us" % (k[:44], v[0], v[1] / 1e3, 100 * v[1] / tot, v[1] / 1e3 / v[0]))
