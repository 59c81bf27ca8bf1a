#!/bin/bash
# multi-GPU bench lines: $1 = number of GPUs, $2 = tag
n=${1:-2}; tag=${2:-multi}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${tag}_bench_${n}gpu.json 2> gpurun_out/${tag}_bench_${n}gpu.err
echo "rc=$?"; tail -3 gpurun_out/${tag}_bench_${n}gpu.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench_${n}gpu.json"))
print("N=%d: %.3f ms/step value %.3e loss %.6f e2e %.3e | grad %.0f us frac %.3f | rowsum %.0f us frac %.3f | clocks %s" % (
    d["n_gpus"], d["ms_per_step"], d["value"], d["loss"], d["e2e"]["value"], d["roofline"]["us_per_launch"], d["roofline"]["frac"],
    d["roofline_rowsum"]["us_per_launch"], d["roofline_rowsum"]["frac"], d["clocks"]))
PY
