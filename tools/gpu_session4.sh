#!/bin/bash
# round 2, fourth GPU session: full suite with the DMMA contraction / sparse host entry / group entry, compute-sanitizer passes,
# and short runs of the other BASELINE configs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider --timeout=900 > gpurun_out/s4_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/s4_pytest.log; grep -E "^FAILED|^ERROR" gpurun_out/s4_pytest.log | head -20
grep -E "f64 .*DMMA|FP64 peak|C4 golden" gpurun_out/s4_pytest.log | cut -c1-220
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s4_memcheck.log 2>&1; echo "memcheck exit $?"; tail -4 gpurun_out/s4_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s4_racecheck.log 2>&1; echo "racecheck exit $?"; tail -4 gpurun_out/s4_racecheck.log
python bench.py --steps 10 --warmup 3 > gpurun_out/s4_bench_c2.json 2> gpurun_out/s4_bench_c2.err; echo "c2 exit $?"
python bench.py --config C1 --steps 2 --warmup 1 > gpurun_out/s4_bench_c1.json 2> gpurun_out/s4_bench_c1.err; echo "c1 exit $?"
python bench.py --config C3 --steps 5 --warmup 3 > gpurun_out/s4_bench_c3.json 2> gpurun_out/s4_bench_c3.err; echo "c3 exit $?"
python bench.py --config C5 --steps 5 --warmup 3 > gpurun_out/s4_bench_c5.json 2> gpurun_out/s4_bench_c5.err; echo "c5 exit $?"
python bench.py --config C4 --steps 2 --warmup 1 --batch 2368 > gpurun_out/s4_bench_c4.json 2> gpurun_out/s4_bench_c4.err; echo "c4 exit $?"
