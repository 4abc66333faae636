#!/bin/bash
# strong-scaling shards on one GPU: the per-GPU batches of N = 2, 4, 8 (what bounds `value` at N GPUs is the slowest rank's step)
mkdir -p gpurun_out
for b in 32768 16384 8192; do
  python bench.py --batch $b --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-exact-extra > gpurun_out/s6_bench_c2_b$b.json 2> gpurun_out/s6_bench_c2_b$b.err
  python - $b <<'P'
import json, sys
b = sys.argv[1]
d = json.load(open(f"gpurun_out/s6_bench_c2_b{b}.json"))
t = d["kernel_timing"]["ms_per_step_instrumented"]
print(b, "ms/step", round(d["ms_per_step"], 3), "step_ms", d["step_ms"], "instrumented", round(t, 3), {k["kernel"][:28]: round(k["share_of_step"] * t, 3) for k in d["kernels"]}, "launches", d["gpu_launches"])
P
done
timeout 300 python -m pytest tests/test_gpu_group.py -m gpu -q -p no:cacheprovider > gpurun_out/s6_pytest_group.log 2>&1; echo "group (1 gpu) exit $?"; tail -2 gpurun_out/s6_pytest_group.log
