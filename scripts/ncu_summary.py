"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one line per kernel in launch order
of the last full step, plus totals."""
import csv, sys, collections
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
seq = []
for r in rows:
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    name = r["Kernel Name"]
    short = name.split("(")[0].replace("void ", "").replace("mdt::", "")[:44]
    seq.append((short, v))
agg = collections.OrderedDict()
for k, v in seq:
    agg.setdefault(k, []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(f"{len(v):4d} x {sum(v)/len(v):8.2f} us  (min {min(v):7.2f})  share {sum(v)/tot:6.1%}  {k}")
# last Adam marks a step end: print the kernels of the final complete step in order
idx = [i for i, (k, _) in enumerate(seq) if k.startswith("adam_flat")]
if len(idx) >= 2:
    a, b = idx[-2] + 1, idx[-1] + 1
    # include the weight-transposes that follow Adam
    step = seq[a + 1: b + 1] if seq[a][0].startswith("transpose") else seq[a:b]
    print("last step:", " | ".join(f"{k[:28]} {v:.1f}" for k, v in step), f" = {sum(v for _, v in step):.1f} us")
