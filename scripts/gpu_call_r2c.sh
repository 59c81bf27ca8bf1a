#!/bin/bash
# round 2, call c: cycle counters of the symmetric row-sum pass (timing build of the library)
mkdir -p gpurun_out
cp muscle_b200/lib/libmuscle_b200.so /tmp/lib_orig.so
cp muscle_b200/lib/exp/libmuscle_b200_0.so muscle_b200/lib/libmuscle_b200.so
MUSCLE_PAIR_TIMING=1 timeout 300 python scripts/pair_ab.py config2 2>&1 | grep -v "^\[pair D" | head -40 | tee gpurun_out/r2c_sym_timing_config2.txt
MUSCLE_PAIR_TIMING=1 timeout 300 python scripts/pair_ab.py config3 2>&1 | grep -A1 "pair_sym" | head -8 | tee gpurun_out/r2c_sym_timing_config3.txt
cp /tmp/lib_orig.so muscle_b200/lib/libmuscle_b200.so
