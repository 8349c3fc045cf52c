"""Print the headline fields of bench.py JSON lines: python tools/show_bench.py gpurun_out/*.json"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e)
        continue
    e, r = d["e2e"], d["roofline"]
    print(f"{f}: N={d['n_gpus']} value {d['value']:.0f} ({d['ms_per_step']:.2f} ms)  e2e {e['value']:.0f} ({e.get('ms_per_step', 0):.2f} ms) "
          f"grp {e.get('group_images')} gather {e.get('gather')}  roof {r['achieved']:.0f} frac {r['frac']:.3f}  sm {d['clocks']['sm_mhz']}")
    for k in ("phases_rank0", "phases_max_over_ranks"):
        if e.get(k):
            print("    ", k, {a: round(v, 2) for a, v in e[k].items()})
    for k in ("parity_vs_single_gpu", "parity_vs_oracle"):
        if e.get(k):
            print("    ", k, ":", e[k])
    if d.get("cpu_baseline"):
        print("     cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["sample"][:90])
