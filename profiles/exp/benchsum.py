import json,sys
d=json.load(open(sys.argv[1]))
print("value", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"] and d["e2e"]["ms_per_step"])
for n,v in (d.get("kernels") or {}).items(): print("%-55s %8.4f ms  %s" % (n, v.get("ms"), ("%.3f" % v["frac_of_hbm_peak"]) if v.get("frac_of_hbm_peak") else ""))
