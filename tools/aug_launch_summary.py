"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of tools/ukf_aug_one.py by kernel.
usage: aug_launch_summary.py launches.csv"""
import collections
import csv
import sys

with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
tot, cnt, seq = collections.OrderedDict(), collections.Counter(), []
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = row["Kernel Name"].split("(")[0]
    if "aug" not in k:
        continue
    v = float(row["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6}.get(row["Metric Unit"], 1.0)
    tot[k] = tot.get(k, 0) + v
    cnt[k] += 1
    seq.append((k, v, row["Grid Size"]))
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    print("%-34s n=%3d %8.3f ms %5.1f%%" % (k, cnt[k], v, 100 * v / s))
print("total %.3f ms over %d launches" % (s, len(seq)))
big = [(k, v, g) for k, v, g in seq if v > 0.5]
print("launches over 0.5 ms:", [("%s %.2f %s" % (k.split("::")[-1][:18], v, g)) for k, v, g in big])
