import json,sys
d=json.loads(sys.stdin.read())
print(round(d["ms_per_step"],2), round(d["ms_per_step_median"],2), [ (s["stage"][:12], round(s["ms"],2)) for s in d["stages"]], d["clocks"]["sm_mhz"], d["clocks"]["reasons"], "e2e", round(d["e2e"]["ms_per_step"],1) if d.get("e2e") else None)
