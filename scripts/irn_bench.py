#!/usr/bin/env python
"""GPU box: the IRN transition matrix (src/indexing.py:116-123, beta = 10, 8 squarings) at hw = 1024 / 4096: three-part
bf16 products on the tensor pipe vs the fp32 GEMM library (MUSCLE_IRN_TC=0)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from muscle_b200 import indexing  # noqa: E402

dev = torch.device("cuda", 0)
for n in (1024, 4096):
    gen = torch.Generator().manual_seed(n)
    a = torch.rand(n, n, generator=gen)
    aff = (0.5 * (a + a.t())).fill_diagonal_(1.0).to(dev)
    rec = dict(hw=n)
    outs = {}
    for mode in ("1", "0"):
        os.environ["MUSCLE_IRN_TC"] = mode
        for _ in range(2):
            outs[mode] = indexing.to_transition_matrix(aff, 10, 8)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            indexing.to_transition_matrix(aff, 10, 8)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        rec["tensor_pipe_ms" if mode == "1" else "library_fp32_ms"] = round(ms, 3)
        rec["tensor_pipe_tflops_fp32_equiv" if mode == "1" else "library_tflops"] = round(8 * 2.0 * n ** 3 / ms / 1e9, 1)
    rec["max_rel_diff"] = float(((outs["1"] - outs["0"]).abs() / outs["0"].abs().clamp_min(1e-30)).max())
    print(json.dumps(rec), file=sys.stderr)
