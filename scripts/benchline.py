import json, sys
lines = [l for l in sys.stdin.read().splitlines() if l.startswith("{")]
d = json.loads(lines[-1])
r = d.get("roofline") or {}
print(sys.argv[1:], "tok/s", round(d["value"], 1), "ms", round(d["ms_per_step"], 4), "frac", round(r.get("frac") or 0, 4),
      "e2e", round(d["e2e"]["value"], 1), "launches", r.get("launches_per_step"), "coll", d.get("collectives_per_step"))
