#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=900 > $O/s20_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/s20_pytest.log; grep -E "^FAILED|^ERROR" $O/s20_pytest.log | head -20
run() { local name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name exit $?"; python tools/bench_brief.py $O/$name.json 2>/dev/null | head -12; }
run r2_bench_c3_v2 --config C3 --no-exact-extra
run r2_bench_c5_v2 --config C5 --no-exact-extra --no-cpu-baseline
run r2_bench_c2_v5 --no-exact-extra --no-cpu-baseline
python tools/single_prof.py
python tools/single_prof.py 504 1216 40 100
