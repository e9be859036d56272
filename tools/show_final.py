"""Print the figures DESIGN.md §6 quotes from a bench.py JSON line (tools/final_runs.sh writes them to gpurun_out/)."""
import json
import sys

r = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("N", r["n_gpus"], "Q1 %.1f G rows/s" % (r["value"] / 1e9), "%.3f ms" % r["ms_per_step"], "plan", r.get("plan"), "kernel frac", r["roofline"].get("frac"),
      "launches", r["gpu_launches"], "clocks", r["clocks"].get("sm_mhz"), r["clocks"].get("reasons"))
print("  e2e %.3f G rows/s" % ((r.get("e2e") or {"value": 0})["value"] / 1e9), "pages ratio", (r.get("e2e_pages") or {}).get("ratio_to_e2e"),
      "plain", ((r.get("e2e_pages") or {}).get("plain_host_pages") or {}).get("ratio_to_e2e"))
for k, v in r["workloads"].items():
    if isinstance(v, dict) and "ms_per_step" in v:
        print("  %-28s %8.3f ms %7.2f G rows/s  frac %s" % (k, v["ms_per_step"], v["rows_per_s"] / 1e9, v.get("hbm_frac_per_gpu")))
q = r["workloads"].get("q3_join_aggregation")
if q:
    print("  q3 stages", q.get("stages_ms_untimed_pass"))
    print("  q3 kernels", q.get("kernels_ms_per_step"))
print("  crosscheck", {k: v for k, v in r.get("crosscheck", {}).items() if isinstance(v, bool)})
