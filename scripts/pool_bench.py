#!/usr/bin/env python
"""GPU box: HBM roofline of the prototype pooling kernels alone (bench.py's pool_record) at config 3 / config 4 sizes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


class _Args:
    gpus = 1


env = bench.Env(_Args())
for n in (65536, 262144):
    r = bench.pool_record(env, n, 128, 21, 30)
    print(json.dumps(dict(pixels=n, fwd_us=round(r["fwd"]["us_per_launch"], 1), fwd_frac=round(r["fwd"]["frac"], 3),
                          bwd_us=round(r["bwd"]["us_per_launch"], 1), bwd_frac=round(r["bwd"]["frac"], 3))), file=sys.stderr)
