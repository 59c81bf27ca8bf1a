#!/bin/bash
# round-2 check aj (2 GPUs): NCCL parity tests and the 2-GPU bench line of the final tree
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py tests/test_gpu_rowsum_sym.py -m gpu -q --timeout 300 2>&1 | tail -6 | tee gpurun_out/r2aj_pytest_nccl_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2aj_bench_2gpu.json 2> gpurun_out/r2aj_bench_2gpu.err
echo "rc=$?"; python - <<PY
import json
d = json.load(open("gpurun_out/r2aj_bench_2gpu.json"))
print("N=%d: %.3f ms/step value %.3e loss %.6f e2e %.3e | grad %.0f us | rowsum %.0f us | clocks %s" % (d["n_gpus"], d["ms_per_step"], d["value"], d["loss"], d["e2e"]["value"], d["roofline"]["us_per_launch"], d["roofline_rowsum"]["us_per_launch"], d["clocks"]))
PY
