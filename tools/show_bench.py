"""Pretty-print the headline numbers of a bench.py JSON line (development helper)."""
import json
import sys

d = json.load(open(sys.argv[1]))
print("value", f"{d['value']:.4g}", "e2e", f"{d['e2e']['value']:.4g}", "roofline frac", round(d["roofline"]["frac"], 4), "mufu",
      round(d["roofline"]["mufu_frac"], 4), d["clocks"], "ms/step", round(d["ms_per_step"], 2))
for k, v in d.get("secondary", {}).items():
    if k == "config5_subset":
        for r in v["shapes"]:
            print("  c5", {kk: (round(vv, 3) if isinstance(vv, float) else vv) for kk, vv in r.items()})
    else:
        print(k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk not in ("workload",) and not kk.endswith("note")})
if "cpu_baseline" in d:
    print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
