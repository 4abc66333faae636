#!/bin/bash
# what the driver runs at round end, in its order: GPU tests, smoke(), the reference arm, the default bench line
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu -p no:cacheprovider > $O/s21_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $O/s21_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/s21_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/s21_smoke.log | cut -c1-160
python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > $O/r2_bench_c2_ref_v2.json 2> $O/r2_bench_c2_ref_v2.err; echo "reference arm exit $?"; cut -c1-300 $O/r2_bench_c2_ref_v2.json
python bench.py > $O/r2_bench_c2_v7.json 2> $O/r2_bench_c2_v7.err; echo "bench exit $?"; python tools/bench_brief.py $O/r2_bench_c2_v7.json
python -c "
import json; d=json.load(open('$O/r2_bench_c2_v7.json')); print({k: d[k] for k in ('metric','value','unit','n_gpus','steps','warmup','ms_per_step','scaling','dtype','gpu_launches')}); print(d['e2e']['value'], d['e2e']['sparse_result']['value'], d['parity'], d['residual']['other_mode']['ms_per_step'])"
