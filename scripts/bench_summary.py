"""One-screen summary of a bench.py JSON line (gpurun tails are short)."""
import json
import sys

for path in sys.argv[1:]:
    try:
        d = json.loads(open(path).read().strip().splitlines()[-1])
    except Exception as e:
        print(path, "unreadable:", e)
        continue
    keep = {k: d.get(k) for k in ("value", "unit", "n_gpus", "ms_per_step", "scaling", "gpu_launches")}
    print(path, keep)
    for k in ("e2e", "roofline", "gemm", "ref_torch_gpu", "cpu_baseline", "clocks", "sharding"):
        if d.get(k) is not None:
            v = d[k]
            if isinstance(v, dict):
                v = {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if not isinstance(b, (dict, list)) or a in ("reasons",)}
            print("  ", k, v)
    for name, rec in sorted((d.get("kernels") or {}).items(), key=lambda kv: -kv[1].get("ms_per_step", 0)):
        print("   kernel %-34s %8.3f ms/step %6.1f launches  %s" % (name, rec.get("ms_per_step", 0), rec.get("launches_per_step", 0),
              ("%.2f of HBM" % rec["frac_of_hbm_peak"]) if "frac_of_hbm_peak" in rec else ""))
