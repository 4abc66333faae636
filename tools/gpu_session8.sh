#!/bin/bash
# the finisher: full suite, then C2 at the full batch and at the strong-scaling shard sizes, with and without it
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider --timeout=900 > gpurun_out/s8_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/s8_pytest.log; grep -E "^FAILED|^ERROR" gpurun_out/s8_pytest.log | head -20
for b in 65536 32768 8192; do
  for fin in 1 0; do
    SSB200_FINISHER=$fin python bench.py --batch $b --steps 20 --warmup 5 --no-cpu-baseline --no-exact-extra > gpurun_out/s8_bench_c2_b${b}_fin$fin.json 2> gpurun_out/s8_bench_c2_b${b}_fin$fin.err
    python - $b $fin <<'P'
import json, sys
b, fin = sys.argv[1], sys.argv[2]
try:
    d = json.load(open(f"gpurun_out/s8_bench_c2_b{b}_fin{fin}.json"))
    t = d["kernel_timing"]["ms_per_step_instrumented"]
    print(b, "finisher", fin, "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), {k["kernel"][:24]: round(k["share_of_step"] * t, 3) for k in d["kernels"]}, "mean iters", d["config"]["mean_iters"])
except Exception as e:
    print(b, fin, "FAILED", e, open(f"gpurun_out/s8_bench_c2_b{b}_fin{fin}.err").read()[-800:])
P
  done
done
