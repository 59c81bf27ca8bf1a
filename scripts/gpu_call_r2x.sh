#!/bin/bash
# round-2 check x: PCM pair passes with HB buffers + two epilogue groups, dW pass with 512-pixel chunks and vector
# reductions, final_kernel with 64-pixel blocks: parity, PCM time + launch list, config-2 launch list, whole suite
mkdir -p gpurun_out
tag=${1:-r2x}
timeout 200 python -m pytest tests/test_gpu_pcm.py -m gpu -q --timeout 60 -x 2>&1 | tail -15 | tee gpurun_out/${tag}_pytest_pcm.log
timeout 120 python scripts/pcm_bench.py 2> gpurun_out/${tag}_pcm_bench.txt; cat gpurun_out/${tag}_pcm_bench.txt
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_pcm_launches.csv \
  python scripts/pcm_bench.py --once > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_pcm_launches.csv | head -12
timeout 300 python bench.py --workload config2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_config2.json 2> gpurun_out/${tag}_config2.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_config2.json"))
print("config2: %.4f ms/step value %.3e e2e %.3e | grad %.1f us frac %.3f | rowsum %.1f us frac %.3f" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["us_per_launch"], d["roofline"]["frac"], d["roofline_rowsum"]["us_per_launch"], d["roofline_rowsum"]["frac"]))
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${tag}_config2_launches.csv \
  python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_config2_launches.csv | head -12
timeout 600 python -m pytest tests -m gpu -q --timeout 120 2>&1 | tail -8 | tee gpurun_out/${tag}_pytest.log
