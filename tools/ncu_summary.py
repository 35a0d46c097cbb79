#!/usr/bin/env python
"""Turn `ncu -i <file>.ncu-rep --page raw --csv` output into the per-kernel JSON summaries kept under profiles/.

usage: ncu -i cap.ncu-rep --page raw --csv > cap.csv; python tools/ncu_summary.py cap.csv out.json "<source note>"
Also writes the DRAM bytes per launch (profiles/top_kernel_traffic.json format) when a fourth argument names it.
"""
import csv
import json
import re
import sys

KEEP = re.compile(r"^(Kernel Name$|dram__bytes_(read|write)\.sum|gpu__dram_throughput\.avg\.pct|gpu__time_duration\.sum$|"
                  r"l1tex__(data_bank_conflicts_pipe_lsu_mem_shared\.sum|t_sector_hit_rate|throughput\.avg\.pct)|"
                  r"launch__(occupancy_limit|registers_per_thread)|lts__t_sector_hit_rate\.pct|"
                  r"sm__inst_executed_pipe_(fp64|tensor_subpipe_dmma|fma|lsu|alu)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__throughput\.avg\.pct|sm__warps_active\.avg\.pct|smsp__average_warps_issue_stalled|"
                  r"smsp__inst_executed\.sum$|smsp__issue_active\.avg\.pct|smsp__pcsamp_warps_issue_stalled)")
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TIME_SCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def main():
    src, dst = sys.argv[1], sys.argv[2]
    note = sys.argv[3] if len(sys.argv) > 3 else ""
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    header, units, data = rows[0], rows[1], rows[2:]
    kernels, traffic = [], {}
    for r in data:
        d = {}
        for name, unit, val in zip(header, units, r):
            if KEEP.match(name):
                d[name] = val if name == "Kernel Name" or not unit else "%s %s" % (val, unit)
        kernels.append(d)
        col = {n: (v, u) for n, u, v in zip(header, units, r)}
        base = re.sub(r"^void |<.*$|\(.*$", "", col["Kernel Name"][0]).split("::")[-1]

        def num(key, scale):
            v, u = col[key]
            return float(v.replace(",", "")) * scale.get(u, 1.0)
        rd, wr = num("dram__bytes_read.sum", UNIT_SCALE), num("dram__bytes_write.sum", UNIT_SCALE)
        traffic[base] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes": rd + wr,
                         "time_ms_under_ncu": num("gpu__time_duration.sum", TIME_SCALE)}
    json.dump({"source": note, "kernels": kernels}, open(dst, "w"), indent=1)
    if len(sys.argv) > 4:
        segs = int(sys.argv[5]) if len(sys.argv) > 5 else 0
        json.dump({"source": dst + " (bench size, per launch)", "segments": segs, "kernels": traffic},
                  open(sys.argv[4], "w"), indent=1)
    for k, v in traffic.items():
        print("%-32s %8.3f ms  dram %.3f GB" % (k, v["time_ms_under_ncu"], v["dram_bytes"] / 1e9))


if __name__ == "__main__":
    main()
