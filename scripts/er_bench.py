#!/usr/bin/env python
"""GPU box: ER loss + the two 8x bilinear up-samplings (csrc/er.cu) fwd+bwd at config 2's full resolution (B=8, 21 x 448 x 448
from stride-8 maps): device ms, L2 flushed between steps.  `--once`: three steps only (for an ncu launch list)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import workloads  # noqa: E402
from muscle_b200.er import er_loss, upsample_bilinear  # noqa: E402

dev = torch.device("cuda", 0)
gen = torch.Generator().manual_seed(0)
cams_s = torch.relu(torch.randn(8, 21, 56, 56, generator=gen)).to(dev)
sgcs_s = torch.relu(torch.randn(8, 21, 56, 56, generator=gen) + 0.3).to(dev).requires_grad_()
lab8 = workloads.synth_labels(8, gen).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def step():
    l = er_loss(upsample_bilinear(cams_s, (448, 448)), upsample_bilinear(sgcs_s, (448, 448)), lab8)
    l.backward()
    sgcs_s.grad = None


for _ in range(3):
    step()
torch.cuda.synchronize()
if "--once" not in sys.argv:
    evs = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); step(); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) for x, y in evs)
    print(json.dumps(dict(er_ms=round(sum(ts) / len(ts), 4), median_ms=round(ts[len(ts) // 2], 4))), file=sys.stderr)
