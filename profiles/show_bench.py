"""Prints the headline fields of a bench.py JSON line: python profiles/show_bench.py FILE..."""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, "N=%d" % d["n_gpus"])
    print("  resnet  %.1f img/s  %.3f ms  e2e %.1f  vs_baseline %s  allreduce_check %s" % (
        d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("vs_baseline"), d.get("allreduce_check")))
    for k, v in (d.get("baselines") or {}).items():
        print("    %-10s %.1f img/s %.3f ms" % (k, v["value"], v["ms_per_step"]))
    w = (d.get("workloads") or {}).get("wide_deep")
    if w:
        print("  w&d     %.2f M/s  %.3f ms  e2e %.2f M/s" % (w["value"] / 1e6, w["ms_per_step"], w["e2e"]["value"] / 1e6))
        for k, v in (w.get("baselines") or {}).items():
            print("    %-10s %.2f M/s %.3f ms ratio %s" % (k, v["value"] / 1e6, v["ms_per_step"], v.get("ratio")))
