"""Print the fields of a bench.py JSON line that the round notes quote.  usage: python profiles/show.py file..."""
import json
import sys

for f in sys.argv[1:]:
    for line in open(f):
        if line.startswith("{"):
            d = json.loads(line)
            e = d.get("e2e", {})
            print(f, d.get("n_gpus"), round(d["value"], 1), d["unit"], round(d["ms_per_step"], 4), "ms",
                  "e2e", round(e.get("value", 0), 1), "launches", d.get("gpu_launches"),
                  "kernel_ms", d.get("roofline", {}).get("kernel_ms"))
