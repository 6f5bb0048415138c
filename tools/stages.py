"""Dev tool: print ms_per_step and the per-stage device times of one bench.py JSON line (stdin)."""
import json
import sys

for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    print("%.3f ms/step | e2e %.3f | " % (d["ms_per_step"], d["e2e"]["ms_per_step"]) +
          " ".join("%s=%.0f" % (k, 1000 * v) for k, v in d["stage_ms"].items()))
