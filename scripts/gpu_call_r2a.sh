#!/bin/bash
# round 2, call a: symmetric row-sum pass — parity first, then A/B timings
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.txt 2>&1
echo "== sym test" ; timeout 600 python -m pytest tests/test_gpu_rowsum_sym.py -q --timeout 300 2>&1 | tail -25 | tee gpurun_out/r2a_pytest_symtest.log
echo "== full suite (sym on)"; timeout 1500 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -25 | tee gpurun_out/r2a_pytest.log
echo "== infonce tests, ordered row-sum pass"; MUSCLE_ROWSUM_SYM=0 timeout 900 python -m pytest tests/test_gpu_infonce.py tests/test_gpu_info_nce.py -q --timeout 600 2>&1 | tail -8 | tee gpurun_out/r2a_pytest_ordered.log
echo "== A/B bg_bias 0.15 (round-1 workload)"
BG_BIAS=0.15 timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | tee gpurun_out/r2a_pair_ab_bias015.txt
BG_BIAS=0.15 MUSCLE_ROWSUM_SYM=0 timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | tee -a gpurun_out/r2a_pair_ab_bias015.txt
echo "== A/B bg_bias 0 (SURVEY 8d workload)"
timeout 300 python scripts/pair_ab.py config2 config3 config4 2>&1 | tee gpurun_out/r2a_pair_ab.txt
MUSCLE_ROWSUM_SYM=0 timeout 300 python scripts/pair_ab.py config2 config3 config4 2>&1 | tee -a gpurun_out/r2a_pair_ab.txt
