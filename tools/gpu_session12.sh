#!/bin/bash
# phase profile of the single-signal kernel (variant build with clock64 timers), cluster vs grid
V=$PWD/build/variants/prof/libssb200.so
python tools/single_prof.py
SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py
SSB200_LIB=$V python tools/single_prof.py
SSB200_LIB=$V SSB200_CLUSTER_SIZE=8 python tools/single_prof.py
SSB200_LIB=$V SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py
SSB200_LIB=$V python tools/single_prof.py 512 2048 32 80
SSB200_LIB=$V SSB200_SINGLE_CLUSTER=0 python tools/single_prof.py 512 2048 32 80
