#!/bin/bash
# bench.py under different env settings, several repetitions:  tools/ab2.sh reps "name|ENV=.." ...
reps=$1; shift
for spec in "$@"; do
  IFS='|' read -r name envs variant <<< "$spec"
  lib=""; [ -n "$variant" ] && lib="SSB200_LIB=$PWD/build/variants/$variant/libssb200.so"
  for r in $(seq $reps); do
    env $envs $lib python bench.py --no-cpu-baseline --no-e2e > gpurun_out/ab2_${name}_$r.json 2> gpurun_out/ab2_${name}_$r.err
    python - "$name" "$r" <<'P'
import json, sys
n, r = sys.argv[1], sys.argv[2]
try:
    d = json.load(open(f"gpurun_out/ab2_{n}_{r}.json")); s = d["step_ms"]
    t = d["kernel_timing"]["ms_per_step_instrumented"]
    ks = {k["kernel"].replace("homotopy_", "")[:22]: (round(k["share_of_step"] * t, 2), round(k.get("frac", 0), 3)) for k in d.get("kernels", [])}
    print(f"{n:10s} {d['ms_per_step']:.2f} ms  median {s['median']:.2f} min {s['min']:.2f} max {s['max']:.2f}  clocks {d['clocks']}  {ks}")
except Exception as ex:
    print(n, "FAILED", ex, open(f"gpurun_out/ab2_{n}_{r}.err").read()[-300:])
P
  done
done
