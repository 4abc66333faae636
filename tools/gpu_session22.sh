#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu -p no:cacheprovider > $O/s22_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $O/s22_pytest.log; grep -E "^FAILED|^ERROR" $O/s22_pytest.log | head
python tests/noise_floor.py --gpu --tag gram24 --modes auto,single > $O/s22_parity.log 2>&1; echo "parity exit $?"
python - <<'P'
import json
d = json.load(open("gpurun_out/parity_gpu_gram24.json"))
for cn, c in d["configs"].items():
    for m in ("auto", "single"):
        if m in c:
            p = c[m]["gpu_vs_ref"]; print(cn[:24].ljust(24), m.ljust(7), [p[x] for x in ("iter", "support", "soft", "coef_1e-4", "coef_1e-3", "err")], p.get("iter_all"))
    p = c["port_vs_ref"]; print(cn[:24].ljust(24), "FLOOR  ", [p[x] for x in ("iter", "support", "soft", "coef_1e-4", "coef_1e-3", "err")], p.get("iter_all"))
P
python tools/single_prof.py
python bench.py > $O/r2_bench_c2_v8.json 2> $O/r2_bench_c2_v8.err; python tools/bench_brief.py $O/r2_bench_c2_v8.json
python bench.py --config C1 --no-cpu-baseline > $O/r2_bench_c1_v6.json 2>/dev/null; python -c "
import json; d=json.load(open('$O/r2_bench_c1_v6.json')); print('C1', round(d['value']), d['config']['us_per_iteration'], d['dropin']['value'])"
