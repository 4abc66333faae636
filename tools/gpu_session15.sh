#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_homotopy.py -m gpu -x -q -p no:cacheprovider --timeout=600 -k "cluster_form or one_call_per_signal or equals_batch_row or full_size or smoke_identity or edge_cases or strided" > $O/s15_single.log 2>&1; echo "single tests exit $?"; tail -5 $O/s15_single.log
python tools/single_prof.py
SSB200_ZERO_COPY=0 python tools/single_prof.py
SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py
python tools/single_prof.py 512 2048 32 80
python tools/single_prof.py 64 256 8 30
SSB200_ZERO_COPY=0 python tools/single_prof.py 64 256 8 30
