#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=900 > $O/s16_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/s16_pytest.log; grep -E "^FAILED|^ERROR" $O/s16_pytest.log | head -20
python tools/single_prof.py
python bench.py --config C1 > $O/r2_bench_c1_v3.json 2> $O/r2_bench_c1_v3.err; python tools/bench_brief.py $O/r2_bench_c1_v3.json
python bench.py --no-exact-extra > $O/r2_bench_c2_v3.json 2> $O/r2_bench_c2_v3.err; python tools/bench_brief.py $O/r2_bench_c2_v3.json
timeout 2400 python bench.py --config GRID > $O/r2_bench_grid_v2.json 2> $O/r2_bench_grid_v2.err; echo "grid exit $?"; tail -3 $O/r2_bench_grid_v2.err
