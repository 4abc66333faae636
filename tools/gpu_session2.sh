#!/bin/bash
# round 2, second GPU session: probes of the coherent (reference bench) workload, re-run of the tests that failed, short C2 bench
mkdir -p gpurun_out
python tests/studies/grid_f64_probe.py --gpu > gpurun_out/s2_grid_f64.log 2>&1; echo "f64 probe exit $?"
python tests/studies/grid_f32_probe.py 256x1024 --max-iters 1024,700 --solves 32 --modes exact,incremental > gpurun_out/s2_grid_f32.log 2>&1; echo "f32 probe exit $?"
SSB200_LIB=$PWD/build/variants/ieee/libssb200.so python tests/studies/grid_f32_probe.py 256x1024 --max-iters 1024 --solves 32 > gpurun_out/s2_grid_f32_ieee.log 2>&1; echo "f32 ieee probe exit $?"
python tests/studies/grid_f32_probe.py 64x256 --max-iters 256,200 --solves 32 > gpurun_out/s2_grid_f32_64x256.log 2>&1
timeout 900 python -m pytest tests/test_gpu_homotopy.py tests/test_gpu_irls.py -m gpu -q -s -p no:cacheprovider -k "one_call or grid_cells or c4 or golden" > gpurun_out/s2_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/s2_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s2_bench_c2.json 2> gpurun_out/s2_bench_c2.err; echo "bench exit $?"
cat gpurun_out/s2_grid_f64.log gpurun_out/s2_grid_f32.log gpurun_out/s2_grid_f32_ieee.log gpurun_out/s2_grid_f32_64x256.log | cut -c1-400
