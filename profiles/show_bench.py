import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1], round(d["value"], 1), "Mcells/s", round(d["ms_per_step"], 3), "ms/step", d["config"]["workload"])
print("  " + "  ".join(f'{k}={v["ms_per_step"]:.3f}' for k, v in d["per_kernel"].items() if v["ms_per_step"] > 0.04))
for k in ("e2e", "file_to_files", "roofline", "roofline_frac_by_kernel", "weti", "slope_gc", "uced", "pit_stat", "relz", "cpu_baseline", "checks", "extra", "clocks"):
    if d.get(k) is not None:
        print(" ", k, json.dumps(d[k])[:900])
