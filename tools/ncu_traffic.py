"""profiles/traffic.json from an `ncu --set full` report: DRAM bytes per launch of our kernels.

usage: ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_traffic.py raw.csv NSYS [out.json]
Also prints a one-line summary per kernel (duration, DRAM GB/s, FP64 pipe, issue, registers) for profiles/*.md.
"""
import csv
import json
import sys

UNITS = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
TIME = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}


def short_name(full):
    n = full.split("(")[0].replace("void ", "").replace("bdilu::", "").replace("wave::", "").replace("quad::", "")
    n = n.replace("(int)", "").replace("(bool)1", "true").replace("(bool)0", "false").replace(" ", "")
    return n


def main(path, nsys, out="profiles/traffic.json"):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name, table=None):
        v = float(r[col[name]].replace(",", ""))
        return v * (table or {}).get(units[col[name]], 1.0)

    kernels = {}
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        name = short_name(r[col["Kernel Name"]])
        rd = val(r, "dram__bytes_read.sum", UNITS)
        wr = val(r, "dram__bytes_write.sum", UNITS)
        dur = val(r, "gpu__time_duration.sum", TIME)
        e = {"nsys": int(nsys), "dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr, "duration_s": dur,
             "dram_gbs": (rd + wr) / dur / 1e9,
             "dram_pct_of_peak": val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
             "fp64_pipe_pct": val(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
             "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "warps_active_pct": val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
             "registers": int(val(r, "launch__registers_per_thread")),
             "grid": r[col["Grid Size"]], "block": r[col["Block Size"]]}
        kernels[name] = e  # last launch of each kernel wins
        print("%-28s grid %-14s %.3f ms  DRAM %.2f GB (r %.2f w %.2f) = %.0f GB/s (%.1f%% of ncu peak)  FP64 pipe %.1f%%  "
              "issue %.1f%%  warps %.1f%%  %d regs" % (name, e["grid"], dur * 1e3, (rd + wr) / 1e9, rd / 1e9, wr / 1e9,
                                                      e["dram_gbs"], e["dram_pct_of_peak"], e["fp64_pipe_pct"],
                                                      e["issue_active_pct"], e["warps_active_pct"], e["registers"]))
    json.dump({"source": path, "kernels": kernels}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], *(sys.argv[3:4]))
