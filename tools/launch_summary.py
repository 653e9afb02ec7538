"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per kernel name+grid count / avg / total."""
import csv, collections, sys
path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.OrderedDict()
total = 0.0
for i, row in enumerate(csv.DictReader(lines)):
    if row.get("Metric Name") != "gpu__time_duration.sum" or i < skip:
        continue
    k = row["Kernel Name"].split("(")[0][-60:] + " grid=" + row["Grid Size"]
    v = float(row["Metric Value"].replace(",", ""))
    agg.setdefault(k, []).append(v)
    total += v
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{len(v):4d} x {sum(v)/len(v)/1e3:9.1f} us = {sum(v)/1e3:10.1f} us ({100*sum(v)/total:5.1f}%)  {k}")
print("total %.1f us" % (total / 1e3))
