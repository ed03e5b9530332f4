import json, sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value", round(d["value"]), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]))
for k in d["roofline"]["full_pass_kernels"]: print("  full %-32s %7.2f us" % (k["kernel"], k["avg_us"]))
