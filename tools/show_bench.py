import json
import sys
d = json.load(open(sys.argv[1]))
print(d["ms_per_step"], d["pcg_iterations"][:6], d["gpu_launches"], "e2e", d["e2e"]["ms_per_step"] if d.get("e2e") else None)
print("roofline", json.dumps(d["roofline"])[:900])
print("pcg", d["pcg_ms_per_iteration"]); print("strong", d.get("strong_1024")); print("cpu", d["cpu_baseline"]); print("parity", d.get("parity"))
for k in d["kernels"]:
    print("%-24s %6.1f %7.3f %7.1f %s" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"], ("%.2f" % k["frac_of_peak"]) if "frac_of_peak" in k else ""))
