import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d["value"]/1e9,3), round(d["ms_per_step"],2), d["roofline"].get("phase_ms"), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
