#!/bin/bash
# round-2 check z: ncu --set full of the PCM pair passes (source-level stall samples)
mkdir -p gpurun_out
tag=${1:-r2z}
timeout 400 ncu --set full --clock-control none --import-source on -k regex:pcm_pair -c 3 -o gpurun_out/${tag}_pcm_pair \
  python scripts/pcm_bench.py --once > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_ncu.log
