"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid, block) count / mean / min / max ms.

usage: python tools/summarize_launches.py launches.csv [title] > profiles/rNN_launches_summary.md
"""
import collections
import csv
import sys


def main(path, title=""):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        name = r["Kernel Name"]
        short = name.split("(")[0][:70]
        rows.append((short, r["Grid Size"], r["Block Size"], ms))
    agg = collections.OrderedDict()
    for k, g, b, ms in rows:
        agg.setdefault((k, g, b), []).append(ms)
    total = sum(ms for *_, ms in rows)
    print("# ncu launch list summary %s" % title)
    print("# ncu --metrics gpu__time_duration.sum --clock-control none ; per-launch times are cold-cache and serialised")
    print("kernel | grid | block | launches | mean ms | min ms | max ms | share of listed time")
    for (k, g, b), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("%s | %s | %s | %d | %.3f | %.3f | %.3f | %.1f%%" % (k, g, b, len(v), sum(v) / len(v), min(v), max(v),
                                                                  100.0 * sum(v) / total))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
