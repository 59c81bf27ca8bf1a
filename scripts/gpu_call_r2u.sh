#!/bin/bash
# round-2 check u (N GPUs): NCCL parity tests, then the bench line with the row-sum pass split by count (default) and
# by pixel block (MUSCLE_ROWSUM_RING=1); $1 = number of GPUs
n=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dist_nccl.py -m gpu -q --timeout 600 2>&1 | tail -8 | tee gpurun_out/r2u_pytest_nccl_${n}gpu.log
for mode in share ring; do
  if [ $mode = ring ]; then export MUSCLE_ROWSUM_RING=1; else unset MUSCLE_ROWSUM_RING; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2u_bench_${n}gpu_${mode}.json 2> gpurun_out/r2u_bench_${n}gpu_${mode}.err
  echo "rc=$?"; tail -2 gpurun_out/r2u_bench_${n}gpu_${mode}.err
  python - <<PY
import json
d = json.load(open("gpurun_out/r2u_bench_${n}gpu_${mode}.json"))
print("${mode} N=%d: %.3f ms/step value %.3e loss %.6f e2e %.3e (%.3f ms) | grad %.0f us | rowsum %.0f us | clocks %s" % (
    d["n_gpus"], d["ms_per_step"], d["value"], d["loss"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["roofline"]["us_per_launch"],
    d["roofline_rowsum"]["us_per_launch"], d["clocks"]))
PY
done
unset MUSCLE_ROWSUM_RING
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 scripts/dist_stage_timing.py --steps 20 > gpurun_out/r2u_stage_timing_${n}gpu.json 2> gpurun_out/r2u_stage_timing_${n}gpu.err
python - <<PY
import json
d = [json.loads(l) for l in open("gpurun_out/r2u_stage_timing_${n}gpu.json") if l.startswith("{")][-1]   # (NCCL prints its banner to stdout)
for k, v in d["stages_ms"].items(): print("%-28s max %.4f min %.4f" % (k, v["max_over_ranks"], v["min_over_ranks"]))
print(d["step_ms"])
PY
