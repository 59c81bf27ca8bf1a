#!/bin/bash
# round-2 check aa: PCM pair-pass variants (buffers, epilogue groups); whole-tile units forward
mkdir -p gpurun_out
tag=${1:-r2aa}
timeout 200 python -m pytest tests/test_gpu_pcm.py -m gpu -q --timeout 60 -x 2>&1 | tail -5 | tee gpurun_out/${tag}_pytest_pcm.log
for fv in 32 31 21; do
  echo "FWD_VAR=$fv" | tee -a gpurun_out/${tag}_pcm_variants.txt
  MUSCLE_PCM_FWD_VAR=$fv timeout 120 python scripts/pcm_bench.py 2>> gpurun_out/${tag}_pcm_variants.txt; tail -1 gpurun_out/${tag}_pcm_variants.txt
done
for bv in 31 22 21; do
  echo "BWD_VAR=$bv" | tee -a gpurun_out/${tag}_pcm_variants.txt
  MUSCLE_PCM_BWD_VAR=$bv timeout 120 python scripts/pcm_bench.py 2>> gpurun_out/${tag}_pcm_variants.txt; tail -1 gpurun_out/${tag}_pcm_variants.txt
done
MUSCLE_PCM_BWD_VAR=21 timeout 200 python -m pytest tests/test_gpu_pcm.py -m gpu -q --timeout 60 -x 2>&1 | tail -3
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_pcm_launches.csv \
  python scripts/pcm_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_pcm_launches.csv 2>/dev/null | head -4
