"""print the parts of a bench.py JSON line that matter when reading a GPU log: python profiles/show_bench.py file.json"""
import json
import sys

d = json.load(open(sys.argv[1]))
print(d["n_gpus"], "GPUs", round(d["ms_per_step"], 2), "ms/step", f"{d['value']:.4g}", "DoF-steps/s, e2e", f"{d['e2e']['value']:.4g}",
      "launches", d["gpu_launches"])
print("iters", d["config"]["gmres_iters_per_step"])
print("parity", d.get("parity"))
print("roofline", d.get("roofline"))
for k, v in (d.get("kernels") or {}).items():
    print("  %-100s" % k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()
                           if a in ("ms", "frac", "share_of_step", "launches_per_step")})
print("step", d.get("step"))
print("clocks", d.get("clocks"))
print("comm", d.get("comm_legs_us"))
print("cpu", (d.get("cpu_baseline") or {}).get("value"))
print("solver", d["config"]["solver"])
