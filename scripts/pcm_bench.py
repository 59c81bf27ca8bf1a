#!/usr/bin/env python
"""GPU box: projection head + PCM (csrc/pcm.cu) fwd+bwd alone at config 2, stride 8 (B=8, 56x56, Cin=336): device ms,
L2 flushed between steps.  `--once` runs three steps only (for an ncu launch list of the same command)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from muscle_b200 import heads  # noqa: E402

dev = torch.device("cuda", 0)
gen = torch.Generator().manual_seed(0)
B, P_h = 8, 56
cam = torch.rand(B, 21, P_h, P_h, generator=gen).to(dev).requires_grad_()
f = torch.relu(torch.randn(B, 336, P_h, P_h, generator=gen)).to(dev)
wt = (torch.randn(128, 336, 1, 1, generator=gen) / 336 ** 0.5).to(dev).requires_grad_()
bias = (torch.randn(128, generator=gen) * 0.1).to(dev).requires_grad_()
go = torch.randn(B, 21, P_h, P_h, generator=gen).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step():
    o = heads.pcm(cam, f, wt, bias)
    o.backward(go)
    cam.grad = wt.grad = bias.grad = None


for _ in range(3):
    step()
torch.cuda.synchronize()
if "--once" not in sys.argv:
    evs = []
    for _ in range(20):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); step(); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) for x, y in evs)
    pairs = B * (P_h * P_h) ** 2
    flops = pairs * 2.0 * (128 + 32) * 2 + pairs * 2.0 * 384
    ms = sum(ts) / len(ts)
    print(json.dumps(dict(pcm_ms=round(ms, 4), median_ms=round(ts[len(ts) // 2], 4), executed_tflops=round(flops / ms / 1e9, 1))),
          file=sys.stderr)
