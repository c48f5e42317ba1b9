import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d["value"]), "MPix/s", round(d["ms_per_step"], 3), "ms/step", {k: round(v["avg_ms"], 3) for k, v in d["kernels"].items()},
      "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"], 3), d["clocks"])
