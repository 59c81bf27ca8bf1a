#!/bin/bash
# round-2 check ah: final tree -- whole GPU suite, smoke(), the default bench line (config 4 + sub-records + pool + ops + CPU baseline)
mkdir -p gpurun_out
tag=${1:-r2ah}
timeout 600 python -m pytest tests -m gpu -q --timeout 120 --durations=5 2>&1 | tail -14 | tee gpurun_out/${tag}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/${tag}_smoke.log
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench.json"))
def line(k, r):
    print("%-8s %.4f ms/step value %.3e e2e %.3e | grad %.1f us %.3f | rowsum %.1f us %.3f | step_tensor %s" % (k, r["ms_per_step"], r["value"], r["e2e"]["value"], r["roofline"]["us_per_launch"], r["roofline"]["frac"], r["roofline_rowsum"]["us_per_launch"], r["roofline_rowsum"]["frac"], r.get("step_tensor")))
line("config4", d)
for k, r in d.get("configs", {}).items(): line(k, r)
p = d["roofline_pool"]
for nm, q in (("pool@c3", p), ("pool@c4", p["at_config4_size"])):
    print(nm, "fwd %.1f us %.3f | bwd %.1f us %.3f" % (q["fwd"]["us_per_launch"], q["fwd"]["frac"], q["bwd"]["us_per_launch"], q["bwd"]["frac"]))
for k, v in d["ops"].items(): print("  op", k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms", "frac_of_bf16_burst", "gbs", "frac_of_hbm")})
print("cpu_baseline", d.get("cpu_baseline"), "clocks", d["clocks"])
PY
