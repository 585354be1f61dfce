"""SASS instruction counts per kernel of the in-tree library -> profiles/rNN_sass_summary.md

usage: python tools/sass_summary.py [rNN] > profiles/rNN_sass_summary.md   (needs cuobjdump + cu++filt from the CUDA toolkit)
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "block_diag_ilu_b200", "libbdilu_b200.so")
COLS = ["UBLKCP", "UTMALDG", "SYNCS", "DFMA", "FFMA", "LDGSTS", "REDUX", "SHFL", "LDS", "STS", "LDG", "STG", "BAR", "MUFU.RCP64H",
        "UTC", "HMMA", "DMMA"]


def main(tag):
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    chunks = re.split(r"\n\s*Function : ", out)[1:]
    rows = []
    for c in chunks:
        name, body = c.split("\n", 1)
        dem = subprocess.run(["cu++filt", name.strip()], capture_output=True, text=True).stdout.strip() or name.strip()
        dem = dem.replace("(int)", "").replace("(bool)1", "true").replace("(bool)0", "false").split("(")[0]
        ops = re.findall(r"/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", body)
        row = [dem, len(ops)]
        for col in COLS:
            if col == "REDUX":
                row.append(sum(1 for o in ops if o.startswith("REDUX") or o.startswith("CREDUX")))
            elif col == "MUFU.RCP64H":
                row.append(sum(1 for o in ops if o.startswith("MUFU.RCP64H")))
            elif col == "UTC":
                row.append(sum(1 for o in ops if o.startswith("UTC")))
            else:
                row.append(sum(1 for o in ops if o.split(".")[0] == col))
        rows.append(row)
    print("# SASS instruction counts per kernel of block_diag_ilu_b200/libbdilu_b200.so (sm_100a), %s\n" % tag)
    print("`cuobjdump -sass` of the in-tree library (tools/sass_summary.py).  UBLKCP = TMA bulk copy (cp.async.bulk), SYNCS = mbarrier,")
    print("LDGSTS = cp.async, REDUX = warp integer reduction (REDUX / CREDUX: pivot search), no UTC*MMA / HMMA / DMMA anywhere")
    print("(north_star: no tensor cores on this path).\n")
    print("| kernel | instrs | " + " | ".join(COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for r in rows:
        print("| `%s` | " % r[0] + " | ".join(str(v) for v in r[1:]) + " |")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "round 2")
