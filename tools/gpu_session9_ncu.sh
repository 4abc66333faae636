#!/bin/bash
# round-2 records: bench lines of every config, the ncu launch list of the C2 step, --set full captures of the dominant kernels.
# The .ncu-rep files are summarised ON THE BOX (tools/ncu_summary.py) and only the text comes back: gpurun_out/ is capped at 64 MiB.
mkdir -p gpurun_out
O=gpurun_out
run() { local name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name exit $?"; python tools/bench_brief.py $O/$name.json 2>/dev/null | head -12; }
run r2_bench_c2_v1
run r2_bench_c1_v1 --config C1
run r2_bench_c3_v1 --config C3
run r2_bench_c5_v1 --config C5
run r2_bench_c4_v1 --config C4
run r2_bench_c2_ref_v1 --impl reference --steps 3 --warmup 1

BENCH="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-exact-extra"
$BENCH > $O/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/ncu_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches.csv $BENCH > $O/ncu_list.log 2>&1
echo "launch list exit $?"
python tools/launch_summary.py $O/r2_launches.csv "$BENCH" > $O/r2_launches_summary.txt; head -14 $O/r2_launches_summary.txt

full() {   # name, skip, count, regex, traffic-config, bench args...
  local name=$1 skip=$2 cnt=$3 rx=$4 tc=$5; shift 5
  timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name "regex:$rx" --launch-skip $skip -c $cnt -f -o $O/$name "$@" > $O/$name.log 2>&1
  echo "$name capture exit $?"
  if [ -n "$tc" ]; then python tools/ncu_summary.py $O/$name.ncu-rep --traffic $O/ncu_traffic.json --config $tc > $O/$name.txt 2>> $O/$name.log
  else python tools/ncu_summary.py $O/$name.ncu-rep > $O/$name.txt 2>> $O/$name.log; fi
  grep -c "^==" $O/$name.txt; ls -la $O/$name.ncu-rep | awk '{print $5}'
  rm -f $O/$name.ncu-rep
}
cp profiles/ncu_traffic.json $O/ncu_traffic.json
K='contract_x3_pair_kernel|homotopy_scan_kernel|homotopy_algebra_kernel'
full r2_ncu_late  90 9 "$K" C2 $BENCH
full r2_ncu_gram   6 6 "$K" C2 $BENCH
full r2_ncu_exact 90 9 "$K" "C2 exact" $BENCH --residual exact
full r2_ncu_dmma   2 2 'contract_dmma_kernel' "" python bench.py --config C4 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
du -sh $O
