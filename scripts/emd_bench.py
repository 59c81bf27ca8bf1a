#!/usr/bin/env python
"""GPU box: crops + EMD dynamic matching (csrc/emd.cu) fwd+bwd at B=8 as in bench_ops: wall ms per step (device time
with the host planning in it) and, with `--once`, three steps for an ncu launch list."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from muscle_b200 import heads  # noqa: E402
from muscle_b200.loss_multilabel import EMD  # noqa: E402
from muscle_b200.torchutils import get_dynamic_crops  # noqa: E402

dev = torch.device("cuda", 0)
gen = torch.Generator().manual_seed(0)
hw = (torch.rand(8, 2, generator=gen) * 100 + 100).long()
c1 = torch.cat([(torch.rand(8, 2, generator=gen) * (224 - hw)).long(), hw], 1)
c2 = torch.cat([(torch.rand(8, 2, generator=gen) * (224 - hw)).long(), hw], 1)
raw1 = F.avg_pool2d(torch.randn(8, 21, 224, 224, generator=gen), 9, 1, 4).to(dev).requires_grad_()
raw2 = F.avg_pool2d(torch.randn(8, 21, 224, 224, generator=gen), 9, 1, 4).to(dev)
v1 = F.normalize(heads.cam_softmaxnorm(raw1), dim=1).detach().requires_grad_()
v2 = F.normalize(heads.cam_softmaxnorm(raw2), dim=1).detach()


def step():
    np.random.seed(3)
    a, b, _ = get_dynamic_crops(v1, c1, v2, c2)
    EMD()(a, b, mode="dynamic").backward()
    v1.grad = None


for _ in range(3):
    step()
torch.cuda.synchronize()
if "--once" not in sys.argv:
    t0 = time.perf_counter()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 10 * 1e3
    t0 = time.perf_counter()
    for _ in range(10):
        np.random.seed(3)
        get_dynamic_crops(v1, c1, v2, c2)
    host = (time.perf_counter() - t0) / 10 * 1e3
    print(json.dumps(dict(emd_step_wall_ms=round(wall, 3), crops_call_host_ms=round(host, 3))), file=sys.stderr)
