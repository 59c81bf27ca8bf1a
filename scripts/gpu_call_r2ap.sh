#!/bin/bash
# round-2 check ap: ER histogram skipping off-label channels, CAM head with coalesced weight staging
mkdir -p gpurun_out
tag=${1:-r2ap}
timeout 300 python -m pytest tests/test_gpu_er.py tests/test_gpu_small.py -m gpu -q --timeout 120 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest.log
timeout 120 python scripts/er_bench.py 2> gpurun_out/${tag}_er_bench.txt; cat gpurun_out/${tag}_er_bench.txt
timeout 120 python scripts/cam_bench.py 2> gpurun_out/${tag}_cam_bench.txt; cat gpurun_out/${tag}_cam_bench.txt
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_er_launches.csv python scripts/er_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_er_launches.csv 2>/dev/null | head -12
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_cam_launches.csv python scripts/cam_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_cam_launches.csv 2>/dev/null | head -8
