"""Summarise an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line:
instructions executed, stall samples, shared-memory wavefronts. Usage: python tools/ncu_lines.py src_both.csv [top]"""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
cur_file, hdr, lines = None, None, []
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
    if r[0] != "" and hdr is not None:      # source-line summary row
        d = dict(zip(hdr[4:], r[4:]))
        def num(k):
            try:
                return float(d.get(k, "0") or 0)
            except ValueError:
                return 0.0
        lines.append((cur_file, int(r[0]), r[1].strip(), num("Instructions Executed"), num("# Samples"),
                      num("Thread Instructions Executed"), num("L1 Wavefronts Shared"), num("Warp Stall Sampling (Not-issued Samples)")))
tot_i = sum(l[3] for l in lines) or 1
tot_s = sum(l[4] for l in lines) or 1
print(f"total warp instructions {tot_i:.3e}, samples {tot_s:.0f}")
print("by instructions executed:")
for l in sorted(lines, key=lambda x: -x[3])[:top]:
    print(f"{l[0][:14]:14s}:{l[1]:4d} inst {100 * l[3] / tot_i:5.1f}% samp {100 * l[4] / tot_s:5.1f}% thr/inst {l[5] / max(l[3], 1):4.1f} | {l[2][:90]}")
print("by stall samples:")
for l in sorted(lines, key=lambda x: -x[4])[:top // 2]:
    print(f"{l[0][:14]:14s}:{l[1]:4d} inst {100 * l[3] / tot_i:5.1f}% samp {100 * l[4] / tot_s:5.1f}% | {l[2][:90]}")
