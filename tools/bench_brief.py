"""One-screen digest of bench.py JSON lines:  python tools/bench_brief.py file.json ..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "ERR", e); continue
    r = d.get("roofline") or {}
    e2e = d.get("e2e") or {}
    cpu = d.get("cpu_baseline") or {}
    print(f"{f}: {d['value']:.0f} {d['unit']}  {d['ms_per_step']:.2f} ms/step  e2e {e2e.get('value', 0):.0f} (h2d {e2e.get('h2d_bytes_per_step', 0) / 1e6:.0f} MB, d2h {e2e.get('d2h_bytes_per_step', 0) / 1e6:.0f} MB)"
          f"  cpu {cpu.get('value', 0):.1f} x{cpu.get('cores')}  clocks {d.get('clocks')}")
    if r:
        print(f"    roofline: {r.get('kernel', '')[:50]} [{str(r.get('bound'))[:8]}] {r.get('achieved', 0):.1f} / {r.get('peak', 0):.1f} {r.get('unit')} = {r.get('frac', 0):.3f}  traffic {r.get('traffic')}")
    for k in d.get("kernels", []):
        print(f"      {k['kernel'][:46]:46s} share {k['share_of_step']:.3f}  frac {k.get('frac', 0):.3f}")
    if "step_ms" in d:
        s = d["step_ms"]
        print(f"    steps: median {s['median']:.2f} min {s['min']:.2f} max {s['max']:.2f}; instrumented {d['kernel_timing']['ms_per_step_instrumented']:.2f}; whole-solve frac {d['whole_solve']['frac']:.3f}")
