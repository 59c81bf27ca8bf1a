#!/bin/bash
# full GPU test suite + a short bench line; $1 = tag for the output files
tag=${1:-suite}
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 --durations=8 2>&1 | tail -40 | tee gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench.json"))
print("config4 N=1: %.2f ms/step value %.3e loss %.6f e2e %.3e | grad %.0f us frac %.3f | rowsum %.0f us frac %.3f | live step frac sustained %.3f" % (
    d["ms_per_step"], d["value"], d["loss"], d["e2e"]["value"], d["roofline"]["us_per_launch"], d["roofline"]["frac"],
    d["roofline_rowsum"]["us_per_launch"], d["roofline_rowsum"]["frac"], d["step_tensor"]["live_frac_of_sustained"]))
for k, v in d.get("configs", {}).items():
    print("%s: %.4f ms/step value %.3e loss %.6f e2e %.3e | grad %.1f us frac %.3f | rowsum %.1f us frac %.3f | live step frac sustained %.3f | cpu %.3e" % (
        k, v["ms_per_step"], v["value"], v["loss"], v["e2e"]["value"], v["roofline"]["us_per_launch"], v["roofline"]["frac"],
        v["roofline_rowsum"]["us_per_launch"], v["roofline_rowsum"]["frac"], v["step_tensor"]["live_frac_of_sustained"], v.get("cpu_baseline", {}).get("value", 0)))
PY
