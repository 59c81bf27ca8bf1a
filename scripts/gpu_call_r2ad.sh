#!/bin/bash
# round-2 check ad: FieldLoss(sobel_size=3), final_kernel with four rows in flight; whole suite; config-2 line + launch list
mkdir -p gpurun_out
tag=${1:-r2ad}
timeout 600 python -m pytest tests -m gpu -q --timeout 120 2>&1 | tail -8 | tee gpurun_out/${tag}_pytest.log
timeout 300 python bench.py --workload config2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_config2.json 2> gpurun_out/${tag}_config2.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_config2.json"))
print("config2: %.4f ms/step value %.3e e2e %.3e | grad %.1f us frac %.3f | rowsum %.1f us frac %.3f" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["us_per_launch"], d["roofline"]["frac"], d["roofline_rowsum"]["us_per_launch"], d["roofline_rowsum"]["frac"]))
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${tag}_config2_launches.csv \
  python bench.py --workload config2 --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python scripts/launch_summary.py gpurun_out/${tag}_config2_launches.csv 2>/dev/null | head -11
