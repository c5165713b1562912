"""Key metrics per kernel out of an `ncu --set full` report:
ncu -i report.ncu-rep --page raw --csv > raw.csv; python tools/ncu_full_summary.py raw.csv "<cmd>" "<what>" """
import csv
import io
import json
import sys

text = open(sys.argv[1], errors="replace").read()
rows = list(csv.DictReader(io.StringIO(text[text.find('"ID"'):])))
WANT = {
    "gpu__time_duration.sum": "time_us",
    "dram__bytes_read.sum": "dram_read_MB",
    "dram__bytes_write.sum": "dram_write_MB",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__shared_mem_per_block_dynamic": "dynamic_smem_bytes",
    "sm__inst_executed.sum": "instructions",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio": "stall_long_scoreboard",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio": "stall_barrier",
    "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio": "stall_short_scoreboard",
    "smsp__average_warp_latency_issue_stalled_mio_throttle.ratio": "stall_mio_throttle",
}
SCALE = {"time_us": 1e-3, "dram_read_MB": 1e-6, "dram_write_MB": 1e-6}
units = rows[0] if rows and rows[0].get("ID", "") == "" else None  # the units row of --page raw
kernels = []
for r in rows:
    if not r.get("ID", "").strip().isdigit():
        continue
    k = {"kernel": r["Kernel Name"].replace("void ", "").replace("unnamed>::", "").replace("rdl::", "")}
    for metric, name in WANT.items():
        v = r.get(metric)
        if v in (None, ""):
            continue
        v = float(v.replace(",", ""))
        unit = (units or {}).get(metric, "")
        if name == "time_us":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
        elif name in ("dram_read_MB", "dram_write_MB"):
            v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1e-6)
        k[name] = v
    kernels.append(k)
print(json.dumps({"command": sys.argv[2] if len(sys.argv) > 2 else "",
                  "what": sys.argv[3] if len(sys.argv) > 3 else "", "kernels": kernels}, indent=1))
