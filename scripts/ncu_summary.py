"""Condense `ncu --page raw --csv` output: one line per captured launch with the metrics the roofline
discussion uses, and dram bytes (read + write) per launch averaged per kernel -> ncu_traffic.json."""
import csv
import json
import re
import sys

src, out_csv, out_json = sys.argv[1:4]
rows = list(csv.reader(open(src)))
hdr = rows[0]
data = rows[2:]          # rows[1] holds the units
want = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "lts__t_bytes.sum",
        "l1tex__t_bytes.sum", "smsp__cycles_active.avg"]
idx = [hdr.index(w) for w in want if w in hdr]
units = rows[1]
with open(out_csv, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([hdr[i] for i in idx])
    w.writerow([units[i] for i in idx])
    for r in data:
        w.writerow([r[i] for i in idx])
traffic = {}
if "dram__bytes_read.sum" in hdr:
    ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in data:
        name = re.sub(r"^void\s+", "", r[ik])
        name = re.split(r"[<(]", name)[0].split("::")[-1]
        b = float(r[ir].replace(",", "")) * scale.get(units[ir], 1) + float(r[iw].replace(",", "")) * scale.get(units[iw], 1)
        acc.setdefault(name, []).append(b)
    for k, v in acc.items():
        traffic[re.sub(r"\d+$", "", k) if k.startswith("k_chol") else k] = sum(v) / len(v)
json.dump(traffic, open(out_json, "w"), indent=1)
print(traffic)
