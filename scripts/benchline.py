import json, sys
d = json.loads(sys.stdin.read())
print(sys.argv[1:], "tok/s", round(d["value"], 1), "ms", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 4), "e2e", round(d["e2e"]["value"], 1))
