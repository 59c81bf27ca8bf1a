#!/bin/bash
# round-2 check ai: ncu of the final tree -- launch list of the bench command, --set full of the PCM and pooling kernels
mkdir -p gpurun_out
tag=${1:-r2ai}
timeout 300 python bench.py --steps 2 --warmup 3 --no-sub --no-cpu-baseline > gpurun_out/${tag}_plain_bench.json 2>/dev/null; echo "plain rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-sub --no-cpu-baseline > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_launches.csv 2>/dev/null | head -12
timeout 400 ncu --set full --clock-control none --import-source on -k regex:pcm_ -c 12 -o gpurun_out/${tag}_pcm \
  python scripts/pcm_bench.py --once > gpurun_out/${tag}_ncu_pcm.log 2>&1; tail -1 gpurun_out/${tag}_ncu_pcm.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:pool_tc -c 6 -o gpurun_out/${tag}_pool \
  python scripts/pool_bench.py > gpurun_out/${tag}_ncu_pool.log 2>&1; tail -1 gpurun_out/${tag}_ncu_pool.log
