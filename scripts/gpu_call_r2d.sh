#!/bin/bash
# round 2, call e: + S from a TMEM-resident row tile
mkdir -p gpurun_out
echo "== sym test" ; timeout 600 python -m pytest tests/test_gpu_rowsum_sym.py tests/test_gpu_infonce.py -q --timeout 300 2>&1 | tail -8 | tee gpurun_out/r2e_pytest_symtest.log
echo "== A/B"
timeout 300 python scripts/pair_ab.py config2 config3 config4 2>&1 | tee gpurun_out/r2e_pair_ab.txt
BG_BIAS=0.15 timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | tee -a gpurun_out/r2e_pair_ab.txt
cp muscle_b200/lib/libmuscle_b200.so /tmp/lib_orig.so
cp muscle_b200/lib/exp/libmuscle_b200_0.so muscle_b200/lib/libmuscle_b200.so
MUSCLE_PAIR_TIMING=1 timeout 300 python scripts/pair_ab.py config2 2>&1 | grep -A1 "pair_sym" | head -4 | tee gpurun_out/r2e_sym_timing_config2.txt
cp /tmp/lib_orig.so muscle_b200/lib/libmuscle_b200.so
