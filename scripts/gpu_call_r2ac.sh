#!/bin/bash
# round-2 check ac: PCM forward pass with one buffer and two CTAs per SM vs three buffers, one CTA per SM
mkdir -p gpurun_out
tag=${1:-r2ac}
timeout 200 python -m pytest tests/test_gpu_pcm.py -m gpu -q --timeout 60 -x 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pcm.log
timeout 120 python scripts/pcm_bench.py 2> gpurun_out/${tag}_pcm_bench.txt; cat gpurun_out/${tag}_pcm_bench.txt
MUSCLE_PCM_FWD_VAR=32 timeout 120 python scripts/pcm_bench.py 2> gpurun_out/${tag}_pcm_bench_32.txt; cat gpurun_out/${tag}_pcm_bench_32.txt
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_pcm_launches.csv \
  python scripts/pcm_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_pcm_launches.csv 2>/dev/null | head -4
