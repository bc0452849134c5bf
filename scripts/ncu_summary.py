"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel (name + grid) the launch count,
average / minimum device time and share of all listed time; then the kernels of the last complete step in launch
order (a step starts at prep_kernel) with each kernel's share of that step."""
import csv, sys, collections
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
seq = []
for r in rows:
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    name = r["Kernel Name"]
    short = name.split("(")[0].replace("void ", "").replace("mdt::", "")[:44]
    if short.startswith(("tc05_gemm_kernel", "latent_", "prep_kernel", "adam_")):
        short += " grid" + r["Grid Size"].replace(" ", "")
    seq.append((short, v))
agg = collections.OrderedDict()
for k, v in seq:
    agg.setdefault(k, []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(f"{len(v):4d} x {sum(v)/len(v):8.2f} us  (min {min(v):7.2f})  share {sum(v)/tot:6.1%}  {k}")
idx = [i for i, (k, _) in enumerate(seq) if k.startswith("prep_kernel")]
if len(idx) >= 2:
    step = seq[idx[-2]: idx[-1]]
    step = [(k, v) for k, v in step if not k.startswith(("<unnamed>", "at::"))]
    st = sum(v for _, v in step)
    print(f"last complete step ({len(step)} kernels, {st:.1f} us serialised, cold cache):")
    for k, v in step:
        print(f"    {v:7.2f} us  {v / st:6.1%}  {k}")
