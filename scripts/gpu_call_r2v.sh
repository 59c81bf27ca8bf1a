#!/bin/bash
# round-2 check v: tensor-pipe pooling (parity on both paths, roofline, one ncu --set full pass), then the whole GPU suite
mkdir -p gpurun_out
tag=r2v
timeout 200 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 60 2>&1 | tail -25 | tee gpurun_out/${tag}_pytest_pool.log
MUSCLE_POOL_TC=0 timeout 200 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 60 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pool_cudacore.log
timeout 120 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench.txt; rc=$?; cat gpurun_out/${tag}_pool_bench.txt
MUSCLE_POOL_TC=0 timeout 120 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench_cudacore.txt; cat gpurun_out/${tag}_pool_bench_cudacore.txt
if [ $rc -eq 0 ]; then
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:pool_ -c 10 -o gpurun_out/${tag}_pool \
    python scripts/pool_bench.py > gpurun_out/${tag}_ncu.log 2>&1
  tail -3 gpurun_out/${tag}_ncu.log
fi
timeout 600 python -m pytest tests -m gpu -q --timeout 120 --durations=5 2>&1 | tail -30 | tee gpurun_out/${tag}_pytest.log
