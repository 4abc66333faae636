#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name "regex:homotopy_single_kernel" --launch-skip 40 -c 1 -f -o $O/r2_ncu_single python tools/single_prof.py > $O/r2_ncu_single.log 2>&1
echo "capture exit $?"
python tools/ncu_summary.py $O/r2_ncu_single.ncu-rep --top 60 > $O/r2_ncu_single.txt 2>> $O/r2_ncu_single.log
# source page with per-line stall samples, trimmed to the hot lines
ncu -i $O/r2_ncu_single.ncu-rep --page source --csv --print-source cuda,sass > $O/r2_ncu_single_source.csv 2>/dev/null
ls -la $O/r2_ncu_single.ncu-rep; rm -f $O/r2_ncu_single.ncu-rep
head -40 $O/r2_ncu_single.txt
