"""Print the headline numbers and the per-kernel table of a bench.py JSON line:  python scripts/bench_summary.py <file.json>"""
import json, sys
d = json.load(open(sys.argv[1]))
print("samples/s %.0f  ms/step %.4f  e2e %.0f  recon %.0f Mpixel/s  launches/step %.1f" %
      (d["value"], d["ms_per_step"], d["e2e"]["value"], d["recon"]["value"], d["gpu_launches"] / d["steps"]))
for k in d["kernels"]:
    print("  %-28s %5.1f/step %7.1f us %5.1f%%" % (k["kernel"], k["launches_per_step"], k["ms_per_launch"] * 1e3, 100 * k["share"]))
for name, c in (d.get("configs") or {}).items():
    print("  %-12s %9.1f samples/s %8.3f ms/step  frac_tf32 %.3f  recon %s" % (name, c["train_samples_per_s"], c["ms_per_step"], c.get("frac_of_tf32_roof", 0),
                                                                         c.get("recon", {}).get("value")))
