#!/bin/bash
# cluster form of the single-signal kernel + one active-set slot per thread in the scan kernel: tests, then A/B benches
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_homotopy.py -m gpu -x -q -p no:cacheprovider --timeout=600 -k "cluster_form or one_call_per_signal or equals_batch_row" > $O/s11_single.log 2>&1; echo "single tests exit $?"; tail -15 $O/s11_single.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=900 > $O/s11_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/s11_pytest.log; grep -E "^FAILED|^ERROR" $O/s11_pytest.log | head -20
brief() { python - "$@" <<'P'
import json, sys
f, tag = sys.argv[1], sys.argv[2]
try:
    d = json.load(open(f))
    print(tag, "value", round(d["value"], 1), d["unit"], "ms/step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1) if d.get("e2e") else None,
          {k: d["config"].get(k) for k in ("us_per_iteration", "mean_iters") if k in d["config"]},
          {k["kernel"][:22]: round(k["share_of_step"] * d["kernel_timing"]["ms_per_step_instrumented"], 3) for k in d.get("kernels", [])} if "kernel_timing" in d else "")
except Exception as e:
    print(tag, "FAILED", e, open(f.replace(".json", ".err")).read()[-600:])
P
}
c1() { local tag=$1; shift; env "$@" python bench.py --config C1 --no-cpu-baseline > $O/s11_c1_$tag.json 2> $O/s11_c1_$tag.err; brief $O/s11_c1_$tag.json "C1 $tag"; }
c1 grid SSB200_SINGLE_CLUSTER=0
c1 cluster16 SSB200_SINGLE_CLUSTER=1
c1 cluster16_nodict SSB200_CLUSTER_DICT=0
c1 cluster8 SSB200_CLUSTER_SIZE=8
c1 cluster4 SSB200_CLUSTER_SIZE=4
c1 v1 SSB200_LIB=$PWD/build/variants/cluster_v1/libssb200.so
c2() { local tag=$1; shift; env "$@" python bench.py --no-cpu-baseline --no-exact-extra > $O/s11_c2_$tag.json 2> $O/s11_c2_$tag.err; brief $O/s11_c2_$tag.json "C2 $tag"; }
c2 kr1 A=1
c2 kr3 SSB200_SCAN_KR3=1
c2 kr1_again A=1
c2 kr3_again SSB200_SCAN_KR3=1
c2 kr1_nocluster SSB200_SINGLE_CLUSTER=0
