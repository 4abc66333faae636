#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_homotopy.py -m gpu -x -q -p no:cacheprovider --timeout=600 -k "cluster_form or one_call_per_signal or equals_batch_row or full_size" > $O/s13_single.log 2>&1; echo "single tests exit $?"; tail -15 $O/s13_single.log
V=$PWD/build/variants/prof/libssb200.so
python tools/single_prof.py
SSB200_CLUSTER_STAGE=0 python tools/single_prof.py
SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py
SSB200_LIB=$V python tools/single_prof.py
SSB200_LIB=$V SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py
python tools/single_prof.py 512 2048 32 80
SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py 512 2048 32 80
python tools/single_prof.py 504 1216 40 100
SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py 504 1216 40 100
