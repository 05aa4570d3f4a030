import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("acq", "%.3e"%d["value"], "e2e %.3e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"], "| trk %.0f"%d["tracking"]["value"], "e2e %.0f"%d["tracking"]["e2e"]["value"], "frac %.3f"%d["tracking"]["roofline"]["frac"])
for k,v in d["roofline"]["kernel_ms"].items(): print("  ",k, "%.4f ms/launch"%(v["total_ms"]/max(1,v["launches"])), v["launches"])
print(d["config"].get("acq_ms_per_step"), d["config"].get("trk_ms_per_step"), d["clocks"])
