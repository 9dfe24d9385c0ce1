"""Per-kernel time shares of the LAST evaluation in an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import re
import sys

path, evals = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 4
rows = list(csv.reader(open(path)))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
data = [(r[kn], float(r[mv].replace(",", ""))) for r in rows[hi + 1:] if len(r) > mv]
per = collections.OrderedDict()
n = len(data) // evals
for k, v in data[-n:]:
    k = re.sub(r"\(.*", "", k)
    k = re.sub(r"void |<unnamed>::", "", k)
    per.setdefault(k, [0, 0])
    per[k][0] += v
    per[k][1] += 1
tot = sum(v[0] for v in per.values())
for k, (v, c) in per.items():
    print(f"{k:50s} {c:3d} launches {v / 1e3:9.1f} us  {100 * v / tot:5.1f}%")
print(f"total {tot / 1e3:.1f} us over {n} launches (ncu per-launch times: cold cache, serialised -- compare shares)")
