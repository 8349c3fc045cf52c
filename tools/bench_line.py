"""tools/bench_line.py — condense bench.py's JSON line (stdin) to the numbers compared between kernel variants."""
import json
import sys

for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    r = d.get("roofline", {})
    print(f"pairs/s {d['value']:.0f}  step {d['ms_per_step']:.1f} ms  kernel {r.get('kernel_ms_per_step', 0):.1f} ms  "
          f"e2e {d.get('e2e', {}).get('value', 0):.0f}  clocks {d.get('clocks', {}).get('sm_mhz')}  matches {d.get('matches_per_step')}")
