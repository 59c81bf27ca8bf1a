#!/bin/bash
# pooling kernels: parity tests (tensor-pipe path and CUDA-core path), then the HBM roofline of both
mkdir -p gpurun_out
tag=${1:-pool}
timeout 240 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 120 -x 2>&1 | tail -15 | tee gpurun_out/${tag}_pytest_pool.log
MUSCLE_POOL_TC=0 timeout 240 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 120 -x 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pool_cudacore.log
timeout 300 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench.txt; cat gpurun_out/${tag}_pool_bench.txt
MUSCLE_POOL_TC=0 timeout 300 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench_cudacore.txt; cat gpurun_out/${tag}_pool_bench_cudacore.txt
