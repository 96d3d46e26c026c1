"""One line per captured launch from the *_summary.txt files scripts/ncu_summary.py writes.
usage: python scripts/ncu_compact.py gpurun_out/r02_ncu_full_*_summary.txt"""
import ast
import sys


def num(s):
    v, _, u = s.partition(" ")
    return float(v.replace(",", "")), u


for path in sys.argv[1:]:
    print("==", path.split("/")[-1].replace("r02_ncu_full_", "").replace("_summary.txt", ""))
    lines = open(path).read().splitlines()
    i = 0
    while i < len(lines):
        if i + 2 < len(lines) and lines[i + 1].lstrip().startswith("{'gpu__time_duration"):
            name = lines[i].strip()
            d = ast.literal_eval(lines[i + 1].strip())
            t = ast.literal_eval(lines[i + 2].split("tensor pipe:", 1)[1].strip())
            dur, du = num(d["gpu__time_duration.sum"])
            dur_ms = dur * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(du, 1.0)
            by = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            rd, ru = num(d["dram__bytes_read.sum"])
            wr, wu = num(d["dram__bytes_write.sum"])
            dram = rd * by[ru] + wr * by[wu]
            tc = t.get("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active", "").split(" ")[0]
            rt = t.get("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "").split(" ")[0]
            x2l = d.get("l1tex__m_xbar2l1tex_read_bytes.sum", "")
            print(f"  {name[:44]:44s} {dur_ms:8.3f} ms  dram {dram / 1e6:9.1f} MB = {dram / dur_ms / 1e6:7.1f} GB/s "
                  f"({d['gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'].split(' ')[0][:5]} %)  L2 "
                  f"{d['lts__throughput.avg.pct_of_peak_sustained_elapsed'].split(' ')[0][:5]} % hit {d['lts__t_sector_hit_rate.pct'].split(' ')[0][:5]} %  "
                  f"L2->SM {x2l}  tensor pipe active {tc[:5]} % (realtime {rt[:5]} %)  regs {d['launch__registers_per_thread'].split(' ')[0]}  "
                  f"grid {d['launch__grid_size']}")
            i += 3
        else:
            i += 1
