#!/bin/bash
# parity record (default build and IEEE-ordering build) and the reference's whole bench grid
mkdir -p gpurun_out
python tests/noise_floor.py --gpu --tag default > gpurun_out/s7_nf_default.log 2>&1; echo "nf exit $?"
SSB200_LIB=$PWD/build/variants/ieee/libssb200.so python tests/noise_floor.py --gpu --tag ieee-order --modes auto,exact > gpurun_out/s7_nf_ieee.log 2>&1; echo "nf ieee exit $?"
timeout 1500 python bench.py --config GRID --batch 32 > gpurun_out/s7_grid.json 2> gpurun_out/s7_grid.err; echo "grid exit $?"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s7_bench_c2.json 2> gpurun_out/s7_bench_c2.err; echo "c2 exit $?"
python bench.py --config C3 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/s7_bench_c3.json 2> gpurun_out/s7_bench_c3.err; echo "c3 exit $?"
python bench.py --config C1 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s7_bench_c1.json 2> gpurun_out/s7_bench_c1.err; echo "c1 exit $?"
