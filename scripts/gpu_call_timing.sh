#!/bin/bash
# cycle counters of the pair passes (timing build of the library in muscle_b200/lib/exp/libmuscle_b200_0.so)
mkdir -p gpurun_out
cp muscle_b200/lib/libmuscle_b200.so /tmp/lib_orig.so
cp muscle_b200/lib/exp/libmuscle_b200_0.so muscle_b200/lib/libmuscle_b200.so
MUSCLE_PAIR_TIMING=1 timeout 300 python scripts/pair_ab.py ${1:-config2} 2>&1 | grep -A3 "${4:-^.pair}" | head -${2:-24} | tee gpurun_out/${3:-timing}.txt
cp /tmp/lib_orig.so muscle_b200/lib/libmuscle_b200.so
