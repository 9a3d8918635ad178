"""Last frame of an ncu launch list (gpu__time_duration.sum, sm__cycles_active.avg, sm__cycles_elapsed.max): one line per kernel."""
import collections
import csv
import re
import sys

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
byid = collections.OrderedDict()
for r in rows:
    d = byid.setdefault(r["ID"], {"k": re.sub(r"\(.*", "", r["Kernel Name"].replace("<unnamed>::", "")), "grid": r["Grid Size"], "stream": r["Stream"]})
    d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
ks = list(byid.values())
last = [i for i, k in enumerate(ks) if k["k"].startswith("k_resolve")]
tot = 0
for k in ks[last[-2] + 1:last[-1] + 1]:
    t = k["gpu__time_duration.sum"] / 1e3
    act = k.get("sm__cycles_active.avg", 0) / max(k.get("sm__cycles_elapsed.max", 1), 1)
    print(f'{k["k"][:48]:48s} s{k["stream"]} grid {k["grid"]:>14s} {t:9.1f} us  sm_active {act:.2f}')
    tot += t
print("sum us", round(tot, 1))
