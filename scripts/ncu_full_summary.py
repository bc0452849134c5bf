"""Summarise an `ncu --set full` report exported with `ncu -i X.ncu-rep --page raw --csv`: one row per kernel with
duration, DRAM bytes, tensor-pipe and occupancy figures."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum"]
idx = {w: hdr.index(w) for w in want if w in hdr}
kn = hdr.index("Kernel Name"); gs = hdr.index("Grid Size")
print("kernel | grid | " + " | ".join(f"{w} [{units[i]}]" for w, i in idx.items()))
for r in data:
    name = r[kn].replace("void ", "").replace("mdt::", "")[:40]
    print(name, "|", r[gs], "|", " | ".join(r[i] for i in idx.values()))
