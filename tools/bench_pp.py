import json,sys
for f in sys.argv[1:]:
    for line in open(f):
        if line.startswith('{"metric"'):
            d=json.loads(line)
            print(f, "Q1 ms", d["ms_per_step"], "frac", d["roofline"]["frac"])
            for k,v in d.get("extra",{}).items():
                if isinstance(v,dict) and "ms_per_step" in v:
                    print("  ",k, v["ms_per_step"],"ms frac",v["hbm_frac_per_gpu"], v["kernels_ms_per_step"])
