#!/bin/bash
# pooling kernels: both test runs, the plain bench, then ONE ncu --set full pass over the pooling kernels of the same command
mkdir -p gpurun_out
tag=${1:-pool}
timeout 240 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 120 -x 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pool.log
MUSCLE_POOL_TC=0 timeout 240 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 120 -x 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pool_cudacore.log
timeout 300 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench.txt; rc=$?; cat gpurun_out/${tag}_pool_bench.txt
if [ $rc -eq 0 ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pool_ -c 12 -o gpurun_out/${tag}_pool \
    python scripts/pool_bench.py > gpurun_out/${tag}_ncu.log 2>&1
  tail -3 gpurun_out/${tag}_ncu.log
fi
