"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the few metrics the roofline uses."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "launch__grid_size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]
idx = [(k, hdr.index(k)) for k in keys if k in hdr]
tens = [(h, i) for i, h in enumerate(hdr) if "tensor" in h and "pct" in h]
for r in rows[2:]:
    print({k: (r[i] + " " + units[i]).strip() for k, i in idx})
    print("   tensor:", {h: r[i] for h, i in tens if r[i] not in ("0", "")})
