"""Host-side cost of one end-to-end step, part by part (config 2).  The pipelined e2e loop of bench.py
is host bound, so this says where the enqueue time goes.  Run on the GPU box:
    python scripts/host_profile.py > gpurun_out/host_profile.txt"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import muscle_b200 as M            # noqa: E402
from muscle_b200 import _lib       # noqa: E402

dev = torch.device("cuda:0")
B, D, H, W, C = 8, 128, 56, 56, 21
x_h = torch.randn(B, D, H, W).pin_memory()
c_h = torch.randint(0, C, (B, H, W), dtype=torch.int64).pin_memory()
copy_stream = torch.cuda.Stream(device=dev)
loss_pinned = torch.empty(4096, dtype=torch.float32).pin_memory()
N = 30      # few enough launches per measurement that the launch queue never pushes back


def timeit(name, fn, n=N, reps=7):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        best = min(best, (t1 - t0) / n)
    print("%-44s %8.1f us/call (host enqueue, best of %d x %d)" % (name, 1e6 * best, reps, n))


def upload(_=0):
    with torch.cuda.stream(copy_stream):
        xg = x_h.to(dev, non_blocking=True)
        cg = c_h.to(dev, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(copy_stream)
    return xg, cg, ev


timeit("upload (2 copies + event on copy stream)", upload)
xg, cg, ev = upload()
torch.cuda.synchronize()
main = torch.cuda.current_stream()


def wait_and_record(_):
    main.wait_event(ev)
    xg.record_stream(main)
    cg.record_stream(main)


timeit("wait_event + 2 record_stream", wait_and_record)
timeit("forward only (no grad)", lambda i: M.dense_infonce(xg, cg, 0.1, C))
xr = xg.clone().requires_grad_()
timeit("forward (grad recorded, never run)", lambda i: M.dense_infonce(xr, cg, 0.1, C))


def fwd_bwd(_):
    xr.grad = None
    M.dense_infonce(xr, cg, 0.1, C).backward()


timeit("forward + backward through autograd", fwd_bwd)

lib = _lib.load()
n = B * H * W
wsb = lib.muscle_dense_infonce_workspace_bytes(n, D, C)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
loss = torch.empty((), device=dev)
dx = torch.empty_like(xg)
one = torch.ones((), device=dev)
sp = main.cuda_stream
xp, cp, lp, dp, wp, op = xg.data_ptr(), cg.data_ptr(), loss.data_ptr(), dx.data_ptr(), ws.data_ptr(), one.data_ptr()
timeit("C ABI muscle_dense_infonce_fwd (ctypes)",
       lambda i: lib.muscle_dense_infonce_fwd(xp, 1, B, H * W, D, cp, 1, C, 0.1, 1e-12, lp, wp, wsb, sp))
timeit("C ABI muscle_dense_infonce_bwd (ctypes)",
       lambda i: lib.muscle_dense_infonce_bwd(1, B, H * W, D, C, 0.1, op, dp, wp, wsb, sp))
timeit("C ABI fwd_bwd (ctypes)",
       lambda i: lib.muscle_dense_infonce_fwd_bwd(xp, 1, B, H * W, D, cp, 1, C, 0.1, 1e-12, lp, dp, wp, wsb, sp))
timeit("torch.empty(ws)", lambda i: torch.empty(wsb, dtype=torch.uint8, device=dev))
timeit("torch.empty(())", lambda i: torch.empty((), device=dev))
timeit("loss_pinned[i].copy_(loss, non_blocking)", lambda i: loss_pinned[i].copy_(loss, non_blocking=True))
timeit("current_stream().cuda_stream", lambda i: torch.cuda.current_stream(dev).cuda_stream)


def whole(i):
    global nxt
    a, b, e = nxt
    main.wait_event(e)
    a.record_stream(main)
    b.record_stream(main)
    nxt = upload()
    a.requires_grad_()
    l = M.dense_infonce(a, b, 0.1, C)
    l.backward()
    loss_pinned[i].copy_(l.detach(), non_blocking=True)


nxt = upload()
timeit("whole pipelined step", whole)


# the same step with two preallocated device buffers instead of allocator traffic per upload
xb = [torch.empty_like(xg) for _ in range(2)]
cb = [torch.empty_like(cg) for _ in range(2)]
up_ev = [torch.cuda.Event() for _ in range(2)]
use_ev = [torch.cuda.Event() for _ in range(2)]
for e in use_ev:
    e.record(main)


def upload2(i):
    b = i & 1
    copy_stream.wait_event(use_ev[b])          # the step that last read this buffer has finished
    with torch.cuda.stream(copy_stream):
        xb[b].copy_(x_h, non_blocking=True)
        cb[b].copy_(c_h, non_blocking=True)
        up_ev[b].record(copy_stream)


def whole2(i):
    b = i & 1
    main.wait_event(up_ev[b])
    upload2(i + 1)
    a = xb[b].detach().requires_grad_()
    l = M.dense_infonce(a, cb[b], 0.1, C)
    l.backward()
    use_ev[b].record(main)
    loss_pinned[i].copy_(l.detach(), non_blocking=True)


upload2(0)
timeit("whole pipelined step, double-buffered", whole2)
