#!/bin/bash
# final evidence of the round: parity counts of the final build (batched and single-call paths), smoke(), the default bench line,
# full-set captures of the late C2 kernels (scan with one slot per thread) and of the single-signal cluster kernel
mkdir -p gpurun_out
O=gpurun_out
python tools/single_prof.py; SSB200_LIB=$PWD/build/variants/prof/libssb200.so python tools/single_prof.py > $O/r2_single_phases.txt; head -3 $O/r2_single_phases.txt
python -m pytest tests/test_gpu_homotopy.py -m gpu -x -q -p no:cacheprovider -k "cluster_form or one_call or out_of_slots or equals_batch" 2>&1 | tail -2
python tests/noise_floor.py --gpu --tag final > $O/s19_parity.log 2>&1; echo "parity exit $?"; tail -3 $O/s19_parity.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > $O/s19_smoke.log 2>&1; echo "smoke exit $?"; tail -4 $O/s19_smoke.log | cut -c1-200
python bench.py > $O/r2_bench_c2_v4.json 2> $O/r2_bench_c2_v4.err; python tools/bench_brief.py $O/r2_bench_c2_v4.json
BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-exact-extra"
full() {   # name, skip, count, regex, traffic-config, command...
  local name=$1 skip=$2 cnt=$3 rx=$4 tc=$5; shift 5
  timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name "regex:$rx" --launch-skip $skip -c $cnt -f -o $O/$name "$@" > $O/$name.log 2>&1
  echo "$name capture exit $?"
  if [ -n "$tc" ]; then python tools/ncu_summary.py $O/$name.ncu-rep --traffic $O/ncu_traffic.json --config "$tc" > $O/$name.txt 2>> $O/$name.log
  else python tools/ncu_summary.py $O/$name.ncu-rep --top 30 > $O/$name.txt 2>> $O/$name.log; fi
  grep -E "^== void" $O/$name.txt | cut -c1-70
  rm -f $O/$name.ncu-rep
}
cp profiles/ncu_traffic.json $O/ncu_traffic.json
K='contract_x3_pair_kernel|homotopy_scan_kernel|homotopy_algebra_kernel'
full r2_ncu_late_v2 46 3 "$K" "C2" $BENCH
full r2_ncu_gram_v2 6 2 "$K" "C2" $BENCH
full r2_ncu_single_v2 40 1 "homotopy_single_kernel" "" python tools/single_prof.py
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches_v2.csv $BENCH > $O/ncu_list.log 2>&1
python tools/launch_summary.py $O/r2_launches_v2.csv "$BENCH" > $O/r2_launches_v2_summary.txt; head -12 $O/r2_launches_v2_summary.txt
du -sh $O
