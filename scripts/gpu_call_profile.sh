#!/bin/bash
# round 2 profiling pass: launch list of the bench command, ncu --set full of the pair passes at config 2 / 3 / 4 and of
# the PCM kernels.  Every ncu command runs only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
set -x
python bench.py --steps 2 --warmup 3 --no-sub --no-cpu-baseline > gpurun_out/r2_plain_bench.json 2> gpurun_out/r2_plain_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-sub --no-cpu-baseline > gpurun_out/r2_ncu_launches.log 2>&1
for wl in config2 config3 config4; do
  python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_plain_$wl.json 2> gpurun_out/r2_plain_$wl.err && \
  ncu --set full --clock-control none --import-source on -k regex:pair_ -s 6 -c 2 -f -o gpurun_out/r2_pair_$wl \
      python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_$wl.log 2>&1
done
python scripts/bench_ops.py --no-cpu --iters 2 > gpurun_out/r2_plain_ops.json 2> gpurun_out/r2_plain_ops.err && \
ncu --set full --clock-control none --import-source on -k regex:pcm_ -s 8 -c 8 -f -o gpurun_out/r2_pcm \
    python scripts/bench_ops.py --no-cpu --iters 2 > gpurun_out/r2_ncu_pcm.log 2>&1
ls -la gpurun_out/*.ncu-rep
