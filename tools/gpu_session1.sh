#!/bin/bash
# round 2, first GPU session: smoke, the GPU test suite, a short C2 bench, the parity record (default and IEEE-ordering builds), GRID cells
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,driver_version --format=csv > gpurun_out/s1_smi.txt 2>&1
nproc >> gpurun_out/s1_smi.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s1_smoke.log 2>&1; echo "smoke exit $?"
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider --timeout=900 > gpurun_out/s1_pytest.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/s1_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/s1_bench_c2.json 2> gpurun_out/s1_bench_c2.err; echo "bench exit $?"
python tests/noise_floor.py --gpu --tag default > gpurun_out/s1_nf_default.log 2>&1; echo "nf exit $?"
SSB200_LIB=$PWD/build/variants/ieee/libssb200.so python tests/noise_floor.py --gpu --tag ieee-order --modes auto,exact > gpurun_out/s1_nf_ieee.log 2>&1; echo "nf ieee exit $?"
timeout 600 python bench.py --config GRID --cells 64x64,64x256,64x2048,256x256,256x1024 --batch 32 > gpurun_out/s1_grid.json 2> gpurun_out/s1_grid.err; echo "grid exit $?"
grep -c PASSED gpurun_out/s1_pytest.log; grep -E "FAILED|ERROR" gpurun_out/s1_pytest.log | head -30
