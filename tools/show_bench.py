import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],4), [(k["kernel"], round(k["us_per_launch"],1), round(k["achieved_gbps"])) for k in d["kernels"]], d["clocks"], d.get("e2e"))
    except Exception as e: print(f, "ERR", e)
