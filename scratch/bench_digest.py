"""Prints the parts of a bench.py JSON line one looks at first."""
import json
import sys

d = json.load(open(sys.argv[1]))
print("ms_per_step", round(d["ms_per_step"], 3), "krylov/step", d["config"]["krylov_iters_per_step"], d["config"].get("linear_solver"),
      "e2e_s", round(d["e2e"]["s_per_step"], 4), "setup_s", round(d["e2e"]["setup_s"], 4))
print("parity", d.get("parity"))
r = d["roofline"]
print("roofline", r["kernel"][:40], "achieved", round(r["achieved"]), "frac", r["frac"] and round(r["frac"], 3), "share", r.get("share_of_step"))
for k, v in d["kernels"].items():
    print("  %-18s ms/step %8.3f  launches %6.1f  GB/s %s" % (k, v["ms_per_step"], v["launches_per_step"], v["GBps"] and round(v["GBps"])))
for k, v in d.get("other_configs", {}).items():
    print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a not in ("kernels", "workload")})
    print("     ", {a: (round(b["ms_per_step"], 3), round(b["GBps"] or 0)) for a, b in v["kernels"].items()})
sc = d.get("same_config", {})
for r in sc.get("rows", []):
    print("  %-62s N=%8d gpu %.5f s  oracle %8.3f s  x%8.1f  krylov %4d  err/tol %.3f" % (
        r["workload"], r["unknowns"], r["gpu_e2e_s"], r["oracle_total_s"], r["speedup_e2e"], r["gpu_krylov_iters"],
        r["max_err_over_tol_vs_oracle"]))
print("oracle growth exponent (3-D):", sc.get("oracle_time_growth_exponent_3d"))
