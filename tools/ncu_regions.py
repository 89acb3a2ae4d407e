"""Region view of an `ncu --page source --print-source cuda,sass --csv` export: SASS instructions in address order, grouped
into runs of equal execution count (basic-block-like), each with the source lines it was attributed to.
Usage: python tools/ncu_regions.py src_both.csv [min_share_percent]"""
import csv
import sys
from collections import Counter

path = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
rows = list(csv.reader(open(path)))
cur_file, hdr, cur_line = None, None, None
sass = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None:
        continue
    if r[0] != "":
        cur_line = int(r[0])
        continue
    d = dict(zip(hdr, r))
    addr = d.get("Address") or r[2]
    try:
        a = int(addr, 16)
    except (ValueError, TypeError):
        continue
    def num(k):
        try:
            return float(d.get(k, "0") or 0)
        except ValueError:
            return 0.0
    sass[a] = (cur_file, cur_line, d.get("Source", "").strip(), num("Instructions Executed"), num("Thread Instructions Executed"), num("# Samples"))
addrs = sorted(sass)
tot = sum(sass[a][3] for a in addrs) or 1
tots = sum(sass[a][5] for a in addrs) or 1
print(f"{len(addrs)} SASS instructions, {tot:.3e} warp instructions executed")
runs = []
for idx, a in enumerate(addrs):
    f, l, src, ie, te, sm = sass[a]
    if runs and abs(runs[-1]["ie"] - ie) <= 0.02 * max(ie, 1):
        rr = runs[-1]
    else:
        rr = {"start": idx, "ie": ie, "n": 0, "sum": 0.0, "thr": 0.0, "samp": 0.0, "lines": Counter(), "ops": Counter()}
        runs.append(rr)
    rr["n"] += 1
    rr["sum"] += ie
    rr["thr"] += te
    rr["samp"] += sm
    rr["lines"][f"{f[:9]}:{l}"] += 1
    rr["ops"][src.split()[0].split(".")[0] if src and not src.startswith("@") else (src.split()[1].split(".")[0] if len(src.split()) > 1 else "?")] += 1
for rr in runs:
    share = 100 * rr["sum"] / tot
    if share < min_share:
        continue
    lines = " ".join(f"{k}x{v}" for k, v in rr["lines"].most_common(6))
    ops = " ".join(f"{k}{v}" for k, v in rr["ops"].most_common(6))
    print(f"@{rr['start']:5d} n={rr['n']:4d} exec/inst={rr['ie']:.3e} share={share:5.2f}% samp={100 * rr['samp'] / tots:5.2f}% thr={rr['thr'] / max(rr['sum'], 1):4.1f} | {lines} | {ops}")
