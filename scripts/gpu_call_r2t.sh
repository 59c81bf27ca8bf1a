#!/bin/bash
# round-2 check t: the allocation-free plan entry alone in a fresh process, the share-split row-sum pass, then the
# whole GPU suite (test_zz_no_alloc.py again, after everything else)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 2>&1 | tail -15 | tee gpurun_out/r2t_pytest_no_alloc_alone.log
timeout 900 python -m pytest tests/test_gpu_rowsum_sym.py -m gpu -q --timeout 600 2>&1 | tail -15 | tee gpurun_out/r2t_pytest_symtest.log
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 --durations=8 2>&1 | tail -40 | tee gpurun_out/r2t_pytest.log
