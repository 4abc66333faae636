#!/bin/bash
# the no-allocation entry alone in a fresh process, then the whole GPU suite (it is collected last), then the two full-set
# captures session 9 aimed wrongly (launch 90 of the three step kernels is the START of the second pass: the late iterations of a
# C2 pass are launches ~31-88)
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > $O/s10_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -3 $O/s10_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout=900 > $O/s10_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/s10_pytest.log; grep -E "^FAILED|^ERROR" $O/s10_pytest.log | head -20
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-exact-extra"
full() {   # name, skip, count, regex, traffic-config, bench args...
  local name=$1 skip=$2 cnt=$3 rx=$4 tc=$5; shift 5
  timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name "regex:$rx" --launch-skip $skip -c $cnt -f -o $O/$name "$@" > $O/$name.log 2>&1
  echo "$name capture exit $?"
  python tools/ncu_summary.py $O/$name.ncu-rep --traffic $O/ncu_traffic.json --config "$tc" > $O/$name.txt 2>> $O/$name.log
  grep -E "^== void" $O/$name.txt | cut -c1-70
  rm -f $O/$name.ncu-rep
}
cp profiles/ncu_traffic.json $O/ncu_traffic.json
K='contract_x3_pair_kernel|homotopy_scan_kernel|homotopy_algebra_kernel'
full r2_ncu_gram2  6 4 "$K" "C2" $BENCH
full r2_ncu_late  46 6 "$K" "C2" $BENCH
full r2_ncu_exact 46 6 "$K" "C2 exact" $BENCH --residual exact
du -sh $O
