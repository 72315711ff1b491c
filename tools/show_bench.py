"""print the headline fields of a bench.py JSON line"""
import json, sys
for l in open(sys.argv[1]):
    if l.startswith("{"):
        d = json.loads(l)
        print("ms", round(d["ms_per_step"], 4), "tok/s", round(d["value"]), "e2e", round((d.get("e2e") or {}).get("value", 0)),
              {k: round(v, 4) for k, v in d.get("stages_ms", {}).items()})
