import json,sys
r=json.load(open(sys.argv[1]))
q=r["workloads"]["q3_join_aggregation"]
print(sys.argv[1], q["ms_per_step"], round(q["rows_per_s"]/1e9,2), all(q["selfcheck"].values())); print(q["kernels_ms_per_step"]); print(q["stages_ms_untimed_pass"])
