#!/bin/bash
# round 2, third GPU session: the separately rounded x update -- reference-bench probes, the whole GPU suite, a short C2 bench
mkdir -p gpurun_out
python tests/studies/grid_f32_probe.py 256x1024 --max-iters 1024 --solves 32 --modes exact,incremental > gpurun_out/s3_grid_f32.log 2>&1; echo "f32 probe exit $?"
python tests/studies/grid_f64_probe.py --gpu > gpurun_out/s3_grid_f64.log 2>&1; echo "f64 probe exit $?"
cat gpurun_out/s3_grid_f32.log gpurun_out/s3_grid_f64.log | cut -c1-330
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider --timeout=900 > gpurun_out/s3_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/s3_pytest.log; grep -E "^FAILED|^ERROR" gpurun_out/s3_pytest.log | head -20
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s3_bench_c2.json 2> gpurun_out/s3_bench_c2.err; echo "bench exit $?"
timeout 900 python bench.py --config GRID --batch 32 --cells 16x256,64x64,64x256,64x1024,64x2048,256x256,256x1024,512x256,512x1024 > gpurun_out/s3_grid.json 2> gpurun_out/s3_grid.err; echo "grid exit $?"
