"""Prints the headline fields of bench.py JSON lines found in the given files (diagnostic helper)."""
import json
import sys

for f in sys.argv[1:]:
    for line in open(f):
        if line.startswith("{"):
            d = json.loads(line)
            st = {k: v["ms"] for k, v in d.get("stages", {}).items()}
            print(f, "n", d.get("n_gpus"), "ms/step", d.get("ms_per_step"), "value", d.get("value"), "e2e",
                  d.get("e2e", {}).get("value"), "clocks", d.get("clocks"), st)
