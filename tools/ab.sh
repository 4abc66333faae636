#!/bin/bash
# A/B of library variants and env knobs on one box:  tools/ab.sh "name|ENV=.. ENV2=..|libvariant" ...
# prints ms/step (per-kernel events on) and each kernel class's ms per step
for spec in "$@"; do
  IFS='|' read -r name envs variant <<< "$spec"
  lib=""
  [ -n "$variant" ] && lib="SSB200_LIB=$PWD/build/variants/$variant/libssb200.so"
  env $envs $lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  env $envs $lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-kernel-timing > gpurun_out/ab_${name}_nt.json 2>> gpurun_out/ab_$name.err
  python - "$name" <<'P'
import json, sys
n = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/ab_{n}.json")); e = json.load(open(f"gpurun_out/ab_{n}_nt.json"))
    ks = {k["kernel"].replace("homotopy_", "")[:28]: round(k["share_of_step"] * d["ms_per_step"], 2) for k in d.get("kernels", [])}
    print(f"{n:10s} timed {d['ms_per_step']:.2f} ms  untimed {e['ms_per_step']:.2f} ms  iters {d['config']['mean_iters']:.4f}  {ks}")
except Exception as ex:
    print(n, "FAILED", ex, open(f"gpurun_out/ab_{n}.err").read()[-400:])
P
done
