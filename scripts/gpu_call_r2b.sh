#!/bin/bash
# round 2, call b: symmetric row-sum pass with non-blocking issuer polls; 3 groups / 2-stage ring vs 2 groups / 3-stage ring; new bench.py
mkdir -p gpurun_out
echo "== sym test (NG=3)" ; timeout 600 python -m pytest tests/test_gpu_rowsum_sym.py -q --timeout 300 2>&1 | tail -8 | tee gpurun_out/r2b_pytest_symtest.log
echo "== sym test (NG=2)" ; MUSCLE_SYM_NG=2 timeout 600 python -m pytest tests/test_gpu_rowsum_sym.py tests/test_gpu_infonce.py -q --timeout 300 2>&1 | tail -8 | tee gpurun_out/r2b_pytest_symtest_ng2.log
echo "== A/B NG=3"
timeout 300 python scripts/pair_ab.py config2 config3 config4 2>&1 | tee gpurun_out/r2b_pair_ab_ng3.txt
BG_BIAS=0.15 timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | tee -a gpurun_out/r2b_pair_ab_ng3.txt
echo "== A/B NG=2"
MUSCLE_SYM_NG=2 timeout 300 python scripts/pair_ab.py config2 config3 config4 2>&1 | tee gpurun_out/r2b_pair_ab_ng2.txt
BG_BIAS=0.15 MUSCLE_SYM_NG=2 timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | tee -a gpurun_out/r2b_pair_ab_ng2.txt
echo "== bench (new contract)"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "rc=$?"; tail -3 gpurun_out/r2b_bench.err; cut -c1-1500 gpurun_out/r2b_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2b_bench_reference.json 2>/dev/null; cut -c1-400 gpurun_out/r2b_bench_reference.json
