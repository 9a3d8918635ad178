"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: the last complete frame's kernels and shares."""
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
names = [(r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", ""), float(r[vi])) for r in rows[1:] if len(r) > vi]
starts = [i for i, (n, _) in enumerate(names) if n.startswith("k_lightgen<0>") or n.startswith("k_lightgen_one")]
frame = names[starts[-2]:starts[-1]] if len(starts) >= 2 else names[starts[-1]:]
tot = sum(v for _, v in frame)
for n, v in frame:
    print(f"{n:48s} {v / 1e3:10.1f} us {100 * v / tot:6.2f} %")
print(f"frame kernels total {tot / 1e6:.3f} ms")
