"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the metrics the roofline uses: duration, DRAM bytes,
L2 traffic / hit rate, tensor-pipe activity, occupancy, registers.  With --json: {kernel: {dram_bytes_per_launch, ...}}.
usage: python scripts/ncu_summary.py file.ncu-rep [--json]"""
import csv
import json
import re
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.max"]
col = {h: i for i, h in enumerate(hdr)}
tensor_cols = [h for h in hdr if ("pipe_tensor" in h or "pipe_tc" in h or "tmem" in h) and ("pct" in h or h.endswith(".sum"))]
summary = {}
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").strip()
    d = {k: (r[col[k]] + " " + units[col[k]]).strip() for k in keys if k in col}
    tens = {h: r[col[h]] + " " + units[col[h]] for h in tensor_cols if r[col[h]] not in ("0", "", "n/a")}
    print(name)
    print("   ", d)
    print("    tensor pipe:", tens)

    def num(k):
        try:
            v = float(r[col[k]].replace(",", ""))
            u = units[col[k]]
            return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1.0, "us": 1e3, "ms": 1e6}.get(u, 1.0)
        except Exception:       # noqa: BLE001
            return None

    s = summary.setdefault(name, dict(per_launch=[], duration_ns=[], tensor_active_pct=[]))
    rd, wr = num("dram__bytes_read.sum"), num("dram__bytes_write.sum")
    if rd is not None and wr is not None:
        s["per_launch"].append(rd + wr)
    s["duration_ns"].append(num("gpu__time_duration.sum"))
    for h in ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"):
        if h in col and r[col[h]] not in ("", "n/a"):
            s["tensor_active_pct"].append(float(r[col[h]].replace(",", "")))
            break
if "--json" in sys.argv:
    for k, s in summary.items():
        if s["per_launch"]:
            s["dram_bytes_per_launch"] = sum(s["per_launch"]) / len(s["per_launch"])
        s["launches_captured"] = len(s["duration_ns"])
    print(json.dumps(summary, indent=1))
