#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=900 > $O/s17_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/s17_pytest.log; grep -E "^FAILED|^ERROR" $O/s17_pytest.log | head -20
timeout 2400 python bench.py --config GRID > $O/r2_bench_grid_v3.json 2> $O/r2_bench_grid_v3.err; echo "grid exit $?"; tail -3 $O/r2_bench_grid_v3.err
python bench.py --config C1 > $O/r2_bench_c1_v4.json 2> $O/r2_bench_c1_v4.err; python tools/bench_brief.py $O/r2_bench_c1_v4.json
