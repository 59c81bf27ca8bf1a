#!/bin/bash
# round-2 check w: pooling with bulk-staged weights + wide finish kernel; config-2 and PCM launch lists
mkdir -p gpurun_out
tag=${1:-r2w}
timeout 200 python -m pytest tests/test_gpu_pool.py -m gpu -q --timeout 60 2>&1 | tail -15 | tee gpurun_out/${tag}_pytest_pool.log
timeout 120 python scripts/pool_bench.py 2> gpurun_out/${tag}_pool_bench.txt; cat gpurun_out/${tag}_pool_bench.txt
timeout 120 python scripts/pcm_bench.py 2> gpurun_out/${tag}_pcm_bench.txt; cat gpurun_out/${tag}_pcm_bench.txt
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_pcm_launches.csv \
  python scripts/pcm_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_pcm_launches.csv | head -20
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${tag}_config2_launches.csv \
  python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_config2_launches.csv | head -20
