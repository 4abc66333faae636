#!/bin/bash
# A/B of library variants and env knobs on one box:  tools/ab.sh "name|ENV=.. ENV2=..|libvariant" ...
# prints ms/step of the timed region, its median step, and each kernel class's ms per step (instrumented pass)
for spec in "$@"; do
  IFS='|' read -r name envs variant <<< "$spec"
  lib=""
  [ -n "$variant" ] && lib="SSB200_LIB=$PWD/build/variants/$variant/libssb200.so"
  env $envs $lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - "$name" <<'P'
import json, sys
n = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/ab_{n}.json"))
    t = d["kernel_timing"]["ms_per_step_instrumented"]
    ks = {k["kernel"].replace("homotopy_", "")[:28]: round(k["share_of_step"] * t, 2) for k in d.get("kernels", [])}
    print(f"{n:10s} {d['ms_per_step']:.2f} ms (median step {d['step_ms']['median']:.2f})  instrumented {t:.2f}  iters {d['config']['mean_iters']:.4f}  {ks}")
except Exception as ex:
    print(n, "FAILED", ex, open(f"gpurun_out/ab_{n}.err").read()[-400:])
P
done
