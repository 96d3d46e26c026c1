"""Summarise an .ncu-rep: headline metrics + top stall-sample SASS lines.  usage: ncu_top.py file.ncu-rep [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__pipe_tensor_cycles_active_realtime.avg.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct", "launch__registers_per_thread", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct", "sm__cycles_elapsed.max.per_second",
        "lts__t_bytes.sum ", "sm__inst_executed_pipe_uniform"]
for k, vals in enumerate(rows[2:]):
    print(f"--- launch {k}: {vals[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else ''}")
    for h, u, v in zip(hdr, units, vals):
        if any(h.startswith(w.strip()) or w in h for w in want) and "TriageCompute" not in h or "pipe_tensor_cycles_active_realtime.avg.pct" in h:
            if not any(x in h for x in (".max.pct", ".min.pct", ".sum.pct_of_peak_sustained_active")):
                print(f"  {h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
hdr = rows[hi]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
sass = [r for r in rows[hi + 1:] if len(r) > isamp and (r[ia].startswith("0x") or r[ia].isdigit())]
tot = sum(int(r[isamp] or 0) for r in sass)
print("total samples", tot)
for r in sorted(sass, key=lambda r: -int(r[isamp] or 0))[:topn]:
    print(f"  {int(r[isamp] or 0):8d} {int(r[iex] or 0):11d}  {r[isrc][:100]}")
