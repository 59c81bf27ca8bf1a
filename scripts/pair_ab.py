#!/usr/bin/env python
"""GPU box: times the pair passes alone (ordered row-sum, symmetric row-sum, gradient) and the whole
single-GPU step on BASELINE config 2 / 3 / 4 shapes.  L2 flushed before every timed launch, CUDA events.
    python scripts/pair_ab.py [config2 config3 ...]      (MUSCLE_ROWSUM_SYM=0 for the ordered step)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import muscle_b200 as M  # noqa: E402
from muscle_b200 import _lib  # noqa: E402
from muscle_b200.stages import CudaStages  # noqa: E402
import workloads  # noqa: E402

SHAPES = {"config2": (8, 128, 56, 56), "config3": (16, 128, 64, 64), "config4": (64, 128, 64, 64),
          "config3_d64": (16, 64, 64, 64)}


def timed(fn, steps, flush):
    evs = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in evs)
    return t[len(t) // 2] * 1e3   # median, us


def main():
    names = sys.argv[1:] or ["config2", "config3"]
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = CudaStages(dev)
    for name in names:
        b, d, h, w = SHAPES[name]
        x, cls, _ = workloads.synth_embeddings(b, d, h, w, 1234, bg_bias=float(os.environ.get("BG_BIAS", "0")))
        xd, cd = x.to(dev), cls.to(dev)
        n = b * h * w
        C = 21
        rp = st.pad_rows(n)
        f32, i32 = torch.float32, torch.int32
        ehat = st.empty((rp, d), torch.bfloat16)
        bits, inv_norm = st.empty((rp,), i32), st.empty((rp,), f32)
        tor, perm, seg = st.empty((rp // 128,), i32), st.empty((n,), i32), st.empty((C + 2,), i32)
        stats = st.zeros((C * d + C,), f32)
        cs, cm = stats[:C * d], stats[C * d:]
        st.prepare(xd, _lib.LAYOUT_NCHW, b, h * w, d, cd.reshape(-1), C, 1e-12, ehat, bits, inv_norm, perm, tor, seg, cs, cm)
        uo, uos = st.empty((rp // 128 + 1,), i32), st.empty((rp // 128 + 1,), i32)
        st.plan(tor, rp, tor, rp, d, uo)
        rowsum, wv, gpos = (st.zeros((rp,), f32) for _ in range(3))
        st.rowsum(ehat, bits, tor, rp, ehat, bits, tor, rp, uo, d, 0.1, rowsum)
        dproto, lacc = st.zeros((C * d,), f32), st.zeros((1,), torch.float64)
        st.mid(ehat, seg, rp, d, C, cs, cm, rowsum, 0.1, wv, gpos, lacc, dproto)
        part = st.empty((st.part_slots(), rp, d), f32)
        tmp = st.zeros((rp,), f32)
        fns = {"rowsum_ordered": lambda: st.rowsum(ehat, bits, tor, rp, ehat, bits, tor, rp, uo, d, 0.1, tmp),
               "grad": lambda: st.grad(ehat, bits, tor, wv, rp, ehat, bits, tor, wv, rp, uo, d, 0.1, part)}
        if d <= 128:
            st.plan_sym(tor, rp, uos)
            fns["rowsum_sym"] = lambda: st.rowsum_sym(ehat, bits, tor, rp, uos, d, 0.1, tmp)

        def step():
            xg = xd.detach().requires_grad_()
            M.dense_infonce(xg, cd, 0.1, C).backward()
        fns["step"] = step
        out = dict(workload=name, n=n, live_units=int(uo[-1]), live_units_sym=int(uos[-1]) if d <= 128 else None,
                   units=(rp // 128) ** 2, sym_env=os.environ.get("MUSCLE_ROWSUM_SYM", "1"))
        for k, fn in fns.items():
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            out[k + "_us"] = round(timed(fn, 20 if n < 100000 else 8, flush), 1)
        live_flops = out["live_units"] * 128.0 * 128.0 * 2 * d
        out["rowsum_ordered_tflops"] = round(live_flops / out["rowsum_ordered_us"] / 1e6, 1)
        if "rowsum_sym_us" in out:
            out["rowsum_sym_tflops_algorithmic"] = round(live_flops / out["rowsum_sym_us"] / 1e6, 1)
        out["grad_tflops"] = round(2 * live_flops / out["grad_us"] / 1e6, 1)
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
