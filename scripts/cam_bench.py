#!/usr/bin/env python
"""GPU box: CAM head + GAP + logits (csrc/small_losses.cu) fwd+bwd alone at config 2, stride 8 (B=8, 56x56, 640 channels):
device ms, L2 flushed between steps.  `--once` runs three steps only (for an ncu launch list of the same command)."""
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
p7 = torch.relu(torch.randn(B, 640, P_h, P_h, generator=gen)).to(dev).requires_grad_()
fw = (torch.randn(21, 640, generator=gen) * 0.05).to(dev).requires_grad_()
gcam = torch.randn(B, 21, P_h, P_h, generator=gen).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step():
    c = heads.cam_head(p7, fw)
    e, l = heads.image_embedding_logits(p7, fw)
    torch.autograd.backward([c, l], [gcam, torch.ones_like(l)])
    p7.grad = fw.grad = None


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
    byt = p7.numel() * 4 * 4 + B * 21 * P_h * P_h * 4 * 3
    ms = sum(ts) / len(ts)
    print(json.dumps(dict(cam_ms=round(ms, 4), median_ms=round(ts[len(ts) // 2], 4), algorithmic_gb_s=round(byt / ms / 1e6, 1))),
          file=sys.stderr)
