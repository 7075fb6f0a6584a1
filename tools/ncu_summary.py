#!/usr/bin/env python
"""Summarise an `ncu --set full` report for profiles/: per kernel (last profiled launch of each name) the duration, DRAM
bytes, hit rates, occupancy and instruction counts, as JSON (read by bench.py for `roofline.traffic`) and as a table.
   python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_ncu_summary"""
import csv
import io
import json
import re
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
header, units, data = rows[0], rows[1], rows[2:]
col = {name: k for k, name in enumerate(header)}
WANT = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "dram__bytes_read.sum.pct_of_peak_sustained_elapsed": "dram_read_pct_of_peak",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "smsp__cycles_active.avg": "cycles_active",
    "sm__cycles_elapsed.max": "cycles_elapsed",
}
SCALE = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0,
         "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def value(row, name):
    if name not in col:
        return None
    txt, unit = row[col[name]].replace(",", ""), units[col[name]]
    try:
        v = float(txt)
    except ValueError:
        return None
    return v * SCALE.get(unit, 1.0)


summary, launches = {}, []
for row in data:
    name = re.sub(r"\(.*", "", row[col["Kernel Name"]]).split("::")[-1].strip()
    entry = {short: value(row, metric) for metric, short in WANT.items()}
    if entry["dram_read"] is not None and entry["dram_write"] is not None:
        entry["dram_bytes_per_launch"] = entry["dram_read"] + entry["dram_write"]
    launches.append((name, entry))
    summary[name] = entry          # the last launch of each kernel wins (warm state)
json.dump(summary, open(out + ".json", "w"), indent=1, sort_keys=True)
with open(out + ".txt", "w") as f:
    f.write(f"# {rep}: ncu --set full --clock-control none; one line per profiled launch\n")
    f.write(f"{'kernel':32s} {'dur_us':>9s} {'dram_rd_MB':>10s} {'dram_wr_MB':>10s} {'dram%':>6s} {'L2hit%':>7s} {'L1hit%':>7s} {'warps%':>7s} {'inst':>11s} {'regs':>5s}\n")
    for name, e in launches:
        g = lambda k, s=1.0: ("%.1f" % (e[k] * s)) if e.get(k) is not None else "-"
        f.write(f"{name[:32]:32s} {g('duration', 1e6):>9s} {g('dram_read', 1e-6):>10s} {g('dram_write', 1e-6):>10s} {g('dram_read_pct_of_peak'):>6s} "
                f"{g('l2_hit_pct'):>7s} {g('l1_hit_pct'):>7s} {g('warps_active_pct'):>7s} {g('warp_instructions'):>11s} {g('registers'):>5s}\n")
print(open(out + ".txt").read())
