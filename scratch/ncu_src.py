"""Summarise `ncu --page source --csv` output: per kernel, instruction mix by opcode and stall samples by reason."""
import csv, sys, collections
path = sys.argv[1]
kern = None; hdr = None
data = collections.OrderedDict()
for row in csv.reader(open(path)):
    if not row: continue
    if row[0] == "Kernel Name":
        kern = row[1][:60]; data[kern] = []; continue
    if row[0] == "Address":
        hdr = row; continue
    if kern and hdr and len(row) == len(hdr):
        data[kern].append(dict(zip(hdr, row)))
for k, rows in data.items():
    print("=" * 100); print(k, len(rows), "SASS instrs")
    tot = sum(int(r["Instructions Executed"]) for r in rows)
    samp = sum(int(r["# Samples"]) for r in rows)
    print("total warp-instr executed", tot, " samples", samp)
    ops = collections.Counter(); ops_s = collections.Counter()
    for r in rows:
        op = r["Source"].split()
        op = [o for o in op if not o.startswith("@")][0].split(".")[0]
        ops[op] += int(r["Instructions Executed"]); ops_s[op] += int(r["# Samples"])
    for op, n in ops.most_common(25):
        print("  %-10s %6.2f%% instr  %6.2f%% samples" % (op, 100.0 * n / tot, 100.0 * ops_s[op] / max(samp, 1)))
    st = collections.Counter()
    for r in rows:
        for h in hdr:
            if h.startswith("stall_") and "Not Issued" not in h:
                st[h] += int(r[h] or 0)
    print("  stalls:", [(a, round(100.0 * b / max(samp, 1), 1)) for a, b in st.most_common(8)])
    if len(sys.argv) > 2:
        top = sorted(rows, key=lambda r: -int(r["# Samples"]))[:int(sys.argv[2])]
        for r in top:
            print("   %5s samp %9s exec  %s" % (r["# Samples"], r["Instructions Executed"], r["Source"][:90]))
