#!/usr/bin/env python
"""Compact view of a bench.py JSON line (headline, e2e, roofline, also entries with their CPU baselines)."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
def f(x): return f"{x:.4g}" if isinstance(x, (int, float)) else str(x)
print("value", f(d["value"]), d["unit"], "| ms/step", f(d["ms_per_step"]), "| n_gpus", d["n_gpus"], "| launches", d.get("gpu_launches"))
if d.get("e2e"): print("e2e", f(d["e2e"]["value"]), "ms/step", f(d["e2e"].get("ms_per_step")))
if d.get("roofline"): r = d["roofline"]; print("roofline", r.get("kernel"), f(r["achieved"]), "GB/s frac", r["frac"], "nominal", r.get("frac_of_nominal_8tbs"), "traffic", r.get("traffic"))
print("breakdown", d.get("step_breakdown")); print("clocks", d.get("clocks"))
for k in ("replicas_identical", "exchange"):
    if k in d: print(k, d[k])
c = d.get("cpu_baseline")
if c: print("cpu_baseline", f(c.get("value")), c.get("kind"), "|", (c.get("sample") or "")[:260])
for e in d.get("also") or []:
    keys = [k for k in e if k.startswith("us_") or k.startswith("ms_")]
    r = e.get("roofline") or {}
    print("--", e["workload"], f(e["value"]), e["unit"], {k: (round(e[k], 2) if isinstance(e[k], float) else e[k]) for k in keys}, "frac", r.get("frac"), "nom", r.get("frac_of_nominal_8tbs"), "traffic", r.get("traffic"))
    for k in ("roofline_sample", "roofline_update", "roofline_one_P1_launch", "gather_roofline"):
        if k in e: print("     ", k, e[k]["frac"], "traffic", e[k].get("traffic"))
    if "bulk_ring_insert" in e: b = e["bulk_ring_insert"]; print("      bulk insert", round(b["us_per_launch"], 1), "us frac", b["roofline"]["frac"])
    c = e.get("cpu_baseline")
    if c: print("      cpu", f(c.get("value")), c.get("kind"), "|", (c.get("sample") or "")[:200])
