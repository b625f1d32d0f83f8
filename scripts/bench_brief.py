"""Print the headline fields of bench.py JSON lines read from stdin (one per line)."""
import json, sys
for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    k = {a: round(b["ms"], 3) for a, b in (d.get("roofline", {}).get("kernels") or {}).items()}
    print(d.get("value"), "tok/s", d.get("ms_per_step"), "ms  e2e", (d.get("e2e") or {}).get("value"), "prefill_ms", d.get("prefill_ms"),
          "roof", d.get("roofline", {}).get("frac"), k, "batched", (d.get("batched") or {}).get("value"))
