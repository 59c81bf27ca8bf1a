#!/usr/bin/env python
"""Benchmark of the contrastive-loss hot path (BASELINE.json metric: dense pixel-pair InfoNCE
fwd+bwd, pixel-pairs/s).

    python bench.py --gpus N --steps K --warmup W          # this repo's CUDA path
    python bench.py --impl reference --gpus N ...          # the reference algorithm on host cores

One "step" = one forward + input-gradient pass of the dense label-masked InfoNCE over one batch
of synthetic EfficientNet-stride-8 embeddings.  N=1: BASELINE config 2 (B=8, 448 crop -> 56x56
pixels, D=128, 21 classes).  N>1: config 4 (global batch 64, 512 crop -> 64x64, D=128) sharded by
image, cross-GPU negatives via all-gather (strong scaling: the global batch is fixed).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "contrastive_loss_fwd_bwd_pixel_pairs_per_s"
UNIT = "pixel-pairs/s"
TEMPERATURE = 0.1
NUM_CLASSES = 21


def workload(n_gpus, name=None):
    if name == "config3":
        return dict(name="config3: dense pseudo-mask pixel-pixel InfoNCE fwd+bwd, batch 16, 512x512 crop, stride 8 "
                         "(64x64), D=128, 21 classes", global_batch=16, h=64, w=64, dim=128)
    if (name is None and n_gpus == 1) or name == "config2":
        return dict(name="config2: MCL dense InfoNCE fwd+bwd, B=8, 448x448 crop, stride 8 (56x56), D=128, 21 classes",
                    global_batch=8, h=56, w=56, dim=128)
    if name == "config5":
        return dict(name="config5 (stress): full pixel-pixel dense InfoNCE with cross-GPU negatives, global batch 256, "
                         "512x512 crop, stride 8 (64x64), D=256, 21 classes",
                    global_batch=256, h=64, w=64, dim=256)
    return dict(name="config4: dense InfoNCE with cross-GPU negatives (all-gather), global batch 64, 512x512 crop, "
                     "stride 8 (64x64), D=128, 21 classes",
                global_batch=64, h=64, w=64, dim=128)


def synth(batch, dim, h, w, seed):
    from oracle import muscle_oracle as O   # synthetic-input generator shared with the tests
    x, cls, _ = O.synth_embeddings(batch, dim, h, w, seed)
    return x, cls


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(bf16_burst=p.get("bf16_tflops"), bf16_sustained=p.get("bf16_tflops_sustained"),
                    hbm=p.get("hbm_gbs"), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index):
        self.index = index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                 "hw_thermal_slowdown": 0x40, "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        s = sorted(self.samples)
        return dict(sm_mhz=(s[len(s) // 2] if s else None), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons),
                    samples=len(s))


def cpu_reference_step(x, cls, q_rows, threads):
    """One fwd+bwd of the reference algorithm (oracle port of src/loss_multilabel.py:9-21 composed
    per SURVEY.md §8c) for the first `q_rows` queries against ALL keys, fp32, on the host cores.
    -> seconds, pixel pairs processed."""
    from oracle import muscle_oracle as O
    torch.set_num_threads(threads)
    d = x.shape[1]
    flat = x.permute(0, 2, 3, 1).reshape(-1, d).clone().requires_grad_()
    c = cls.reshape(-1)
    t0 = time.perf_counter()
    e = O.l2_normalize_pixels(flat)
    proto, _ = O.class_prototypes(e, c, NUM_CLASSES)
    loss = O.dense_infonce(e[:q_rows], c[:q_rows], e, c, TEMPERATURE, NUM_CLASSES, proto=proto, row_chunk=2048)
    loss.backward()
    dt = time.perf_counter() - t0
    return dt, float(q_rows) * flat.shape[0]


def run_reference(args, wl):
    """--impl reference: the reference algorithm timed on this box's host cores (bounded sample)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    x, cls = synth(wl["global_batch"], wl["dim"], wl["h"], wl["w"], 1234)
    n = x.shape[0] * x.shape[2] * x.shape[3]
    q_rows = min(n, 4096)
    for _ in range(min(args.warmup, 1)):
        cpu_reference_step(x, cls, q_rows, threads)
    steps = max(1, min(args.steps, 5))
    tot_t, tot_p = 0.0, 0.0
    for _ in range(steps):
        dt, pairs = cpu_reference_step(x, cls, q_rows, threads)
        tot_t += dt
        tot_p += pairs
    v = tot_p / tot_t
    sample = "%d query pixels x all %d key pixels per step, %d steps, fp32, torch CPU" % (q_rows, n, steps)
    line = dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=steps,
                warmup=min(args.warmup, 1), ms_per_step=1e3 * tot_t / steps, higher_is_better=True,
                scaling="strong" if args.gpus > 1 else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["name"]),
                cpu_baseline=dict(value=v, unit=UNIT, cores=threads, kind="port", sample=sample),
                e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default=None, choices=["config2", "config3", "config4", "config5"],
                    help="override the default (config2 at 1 GPU, config4 at N>1), e.g. config4 on 1 GPU")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = workload(args.gpus, args.workload)
    if args.impl == "reference":
        return run_reference(args, wl)

    import torch.distributed as dist
    import muscle_b200 as M
    from muscle_b200 import dist as MD
    from muscle_b200.stages import CudaStages
    from muscle_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL prints its version banner on STDOUT at NCCL_DEBUG=VERSION; stdout carries the one JSON line
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    gb, h, w, dim = wl["global_batch"], wl["h"], wl["w"], wl["dim"]
    assert gb % world == 0
    lb = gb // world
    x_all, cls_all = synth(gb, dim, h, w, 1234)
    x_h = x_all[rank * lb:(rank + 1) * lb].contiguous().pin_memory()
    c_h = cls_all[rank * lb:(rank + 1) * lb].contiguous().pin_memory()
    x_d = x_h.to(dev)
    c_d = c_h.to(dev)
    n_local = lb * h * w
    n_global = gb * h * w
    pairs_per_step = float(n_global) * float(n_global)     # whole job: every (query, key) entry of the global matrix

    def step_device():
        xg = x_d.detach().requires_grad_()
        if world > 1:
            loss = MD.dense_infonce_sharded(xg, c_d, TEMPERATURE, NUM_CLASSES)
        else:
            loss = M.dense_infonce(xg, c_d, TEMPERATURE, NUM_CLASSES)
        loss.backward()
        return loss, xg.grad

    def step_e2e():
        xg = x_h.to(dev, non_blocking=True).requires_grad_()
        cg = c_h.to(dev, non_blocking=True)
        if world > 1:
            loss = MD.dense_infonce_sharded(xg, cg, TEMPERATURE, NUM_CLASSES)
        else:
            loss = M.dense_infonce(xg, cg, TEMPERATURE, NUM_CLASSES)
        loss.backward()
        return float(loss.item())          # device -> host read of the step's result

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """Per-step CUDA-event timing on the launching stream, L2 flushed (untimed) before each step."""
        evs = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / 1e3

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_dev = timed(step_device, args.steps)
    barrier()
    clocks = sampler.stop()
    for _ in range(3):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    t_e2e_serial = time.perf_counter() - t0
    barrier()

    # The same through the public API the way a training loop feeds it: every step's inputs still go
    # pinned host -> device and every step's loss still comes back to the host inside the timed region,
    # but the next step's upload runs on a copy stream under the current step's kernels and the loss is
    # read back asynchronously (checked one step later), so no step waits on PCIe or on a host sync.
    copy_stream = torch.cuda.Stream(device=dev)
    loss_pinned = torch.empty(args.steps + 3, dtype=torch.float32).pin_memory()

    # two device-resident input buffers filled in turn (the usual prefetcher): no allocator traffic per step
    x_buf = [torch.empty(x_h.shape, dtype=x_h.dtype, device=dev) for _ in range(2)]
    c_buf = [torch.empty(c_h.shape, dtype=c_h.dtype, device=dev) for _ in range(2)]
    up_ev = [torch.cuda.Event() for _ in range(2)]     # upload into buffer b has finished
    use_ev = [torch.cuda.Event() for _ in range(2)]    # the step that read buffer b has finished

    def upload(i):
        b = i & 1
        copy_stream.wait_event(use_ev[b])
        with torch.cuda.stream(copy_stream):
            x_buf[b].copy_(x_h, non_blocking=True)
            c_buf[b].copy_(c_h, non_blocking=True)
            up_ev[b].record(copy_stream)

    def e2e_pipelined(steps, fused=False):
        main = torch.cuda.current_stream()
        for e in use_ev:
            e.record(main)
        gc_was = gc.isenabled()
        gc.disable()                  # a collection in the middle of a 0.3 ms step costs several steps
        t_host0 = time.perf_counter()
        upload(0)
        for i in range(steps):
            b = i & 1
            main.wait_event(up_ev[b])
            if i + 1 < steps:
                upload(i + 1)
            if fused:       # loss and dx from one C call, no autograd (dense_infonce_value_and_grad)
                loss, _dx = M.dense_infonce_value_and_grad(x_buf[b], c_buf[b], TEMPERATURE, NUM_CLASSES)
            else:
                xg = x_buf[b].detach().requires_grad_()
                if world > 1:
                    loss = MD.dense_infonce_sharded(xg, c_buf[b], TEMPERATURE, NUM_CLASSES)
                else:
                    loss = M.dense_infonce(xg, c_buf[b], TEMPERATURE, NUM_CLASSES)
                loss.backward()
            use_ev[b].record(main)
            loss_pinned[i].copy_(loss.detach(), non_blocking=True)   # device -> host read of the step's result
        e2e_pipelined.host_s = time.perf_counter() - t_host0          # all steps enqueued (host side only)
        torch.cuda.synchronize()
        if gc_was:
            gc.enable()
        return loss_pinned[:steps].tolist()

    e2e_pipelined(3)
    # the loop is host bound by a small margin (enqueue ~0.27 ms/step against ~0.28 ms of kernels at
    # config 2), so one descheduled host thread doubles a run: three runs of K steps, the median reported
    e2e_runs = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        e2e_losses = e2e_pipelined(args.steps)
        e2e_runs.append((time.perf_counter() - t0, e2e_pipelined.host_s))
        assert all(v == v for v in e2e_losses), "NaN loss in the end-to-end run"
    e2e_runs.sort()
    t_e2e, e2e_host_s = e2e_runs[1]
    # the uploads alone (same buffers, same stream): the PCIe floor of any end-to-end step
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        upload(i)
    torch.cuda.synchronize()
    t_h2d = time.perf_counter() - t0
    fused_e2e = None
    if world == 1:      # informational: the same loop through the autograd-free public call
        e2e_pipelined(3, fused=True)
        fr = []
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            e2e_pipelined(args.steps, fused=True)
            fr.append((time.perf_counter() - t0, e2e_pipelined.host_s))
        fr.sort()
        fused_e2e = dict(ms_per_step=1e3 * fr[1][0] / args.steps, value=pairs_per_step * args.steps / fr[1][0],
                         host_enqueue_ms_per_step=1e3 * fr[1][1] / args.steps,
                         note="dense_infonce_value_and_grad: loss and dx from one C call, no autograd engine")
    barrier()

    # dominant kernel (gradient pass) and the row-sum pass, timed alone through the staged C ABI
    st = CudaStages(dev)
    rows_pad = st.pad_rows(n_local)
    kern = {}
    if world == 1:
        cols_pad = rows_pad
        f32, i32 = torch.float32, torch.int32
        C = NUM_CLASSES
        ehat = st.empty((rows_pad, dim), torch.bfloat16)
        bits = st.empty((rows_pad,), i32)
        tor = st.empty((rows_pad // 128,), i32)
        inv_norm = st.empty((rows_pad,), f32)
        perm = st.empty((n_local,), i32)
        seg = st.empty((C + 2,), i32)
        stats = st.zeros((C * dim + C,), f32)
        cs, cm = stats[:C * dim], stats[C * dim:]
        rowsum, wv, gpos = (st.zeros((rows_pad,), f32) for _ in range(3))
        dproto = st.zeros((C * dim,), f32)
        lacc = st.zeros((1,), torch.float64)
        st.prepare(x_d, _lib.LAYOUT_NCHW, lb, h * w, dim, c_d, C, 1e-12, ehat, bits, inv_norm, perm, tor, seg, cs, cm)
        unit_off = st.empty((rows_pad // st.unit_rows(dim) + 1,), i32)
        st.plan(tor, rows_pad, tor, cols_pad, dim, unit_off)
        st.rowsum(ehat, bits, tor, rows_pad, ehat, bits, tor, cols_pad, unit_off, dim, TEMPERATURE, rowsum)
        st.mid(ehat, seg, rows_pad, dim, C, cs, cm, rowsum, TEMPERATURE, wv, gpos, lacc, dproto)
        part = st.empty((st.part_slots(), rows_pad, dim), f32)
        grad_fn = lambda: st.grad(ehat, bits, tor, wv, rows_pad, ehat, bits, tor, wv, cols_pad, unit_off, dim, TEMPERATURE, part)
        rs_tmp = st.zeros((rows_pad,), f32)
        rowsum_fn = lambda: st.rowsum(ehat, bits, tor, rows_pad, ehat, bits, tor, cols_pad, unit_off, dim, TEMPERATURE, rs_tmp)
        live_units = int(unit_off[-1].item())
        for _ in range(3):
            grad_fn(); rowsum_fn()
        ksteps = max(10, min(args.steps, 50))
        kern["grad_s"] = timed(grad_fn, ksteps) / ksteps
        kern["rowsum_s"] = timed(rowsum_fn, ksteps) / ksteps
        kern["live_frac"] = live_units / float((rows_pad // st.unit_rows(dim)) * (cols_pad // 128))

    # max over ranks (device-timed)
    t = torch.tensor([t_dev, t_e2e, t_e2e_serial], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t_dev, t_e2e, t_e2e_serial = float(t[0]), float(t[1]), float(t[2])

    if rank == 0:
        pk = peaks()
        value = pairs_per_step * args.steps / t_dev
        e2e_v = pairs_per_step * args.steps / t_e2e
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=1e3 * t_dev / args.steps, higher_is_better=True,
                    scaling="strong" if world > 1 else "weak", vs_baseline=None, dtype="bf16", data="synthetic",
                    config=dict(workload=wl["name"], global_batch=gb, pixels_per_image=h * w, proj_dim=dim,
                                temperature=TEMPERATURE, parallelism="dp%d (batch-sharded, all-gather keys)" % world,
                                l2="flushed (256 MiB write) before every timed step",
                                accumulate="fp32 (TMEM); operands bf16"),
                    e2e=dict(value=e2e_v, unit=UNIT, h2d_bytes_per_step=int(x_h.numel() * 4 + c_h.numel() * 8),
                             d2h_bytes_per_step=4, ms_per_step=1e3 * t_e2e / args.steps,
                             mode="pipelined: next step's pinned-host upload on a copy stream into one of two device buffers, loss read back "
                                  "asynchronously; every step's H2D and D2H are inside the timed region",
                             host_enqueue_ms_per_step=1e3 * e2e_host_s / args.steps,
                             value_and_grad=fused_e2e,
                             h2d_only_ms_per_step=1e3 * t_h2d / args.steps,
                             h2d_gb_per_s=(x_h.numel() * 4 + c_h.numel() * 8) * args.steps / t_h2d / 1e9,
                             runs_ms_per_step=[1e3 * r[0] / args.steps for r in e2e_runs],
                             runs_note="three runs of `steps` steps each, the median is `value`",
                             serial_value=pairs_per_step * args.steps / t_e2e_serial,
                             serial_ms_per_step=1e3 * t_e2e_serial / args.steps,
                             serial_note="upload -> step -> blocking .item() per step, nothing overlapped"),
                    gpu_launches=int(args.steps * 8), clocks=clocks)
        flops_step = 6.0 * dim * pairs_per_step
        line["step_tensor"] = dict(algorithmic_tflops_per_gpu=flops_step * args.steps / t_dev / 1e12 / world,
                                   frac_of_sustained=flops_step * args.steps / t_dev / 1e12 / world / pk["bf16_sustained"],
                                   note="per GPU; 6*D flops per ordered pixel pair (S, dQ, dK) over ALL pairs, whole step "
                                        "incl. HBM kernels and collectives; tile pairs whose pixels share one class are "
                                        "fully masked and skipped, so this can exceed the executed tensor rate")
        if kern:
            # algorithmic flops of a launch = the LIVE 128 x 128 tile pairs only: pairs whose tiles hold one and
            # the same class are masked by the loss and are never multiplied
            allp = 4.0 * dim * float(n_local) * float(n_global) / kern["grad_s"] / 1e12
            ach = allp * kern["live_frac"]
            # DRAM bytes per launch of this kernel from the committed ncu --set full capture (same workload only)
            traffic, traffic_src = None, None
            tpath = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tpath) and wl["name"].startswith("config2") and world == 1:
                with open(tpath) as f:
                    tj = json.load(f)
                traffic = tj["bytes_per_launch"].get("pair_kernel<%d,MODE_GRAD>" % dim)
                traffic_src = tj["source"]
            line["roofline"] = dict(bound="tensor", kernel="pair_kernel<%d,MODE_GRAD>" % dim, achieved=ach,
                                    peak=pk["bf16_burst"], unit="TFLOP/s", frac=ach / pk["bf16_burst"],
                                    frac_of_sustained=ach / pk["bf16_sustained"], traffic=traffic,
                                    traffic_source=traffic_src,
                                    traffic_note="algorithmic DRAM bytes per launch = the bf16 operand once: "
                                                 "%d bytes" % (rows_pad * dim * 2),
                                    peak_source=pk["source"] + ", burst figure (kernel timed alone)",
                                    algorithmic="4*D flops per ordered pixel pair of every live (row tile, column tile) "
                                                "unit: S recomputed (2*D) + one T@E product (2*D) that serves dQ and dK "
                                                "through the symmetric form",
                                    us_per_launch=1e6 * kern["grad_s"], live_tile_fraction=kern["live_frac"],
                                    all_pairs_tflops=allp,
                                    all_pairs_note="the same launch counted over ALL pixel pairs, masked ones included "
                                                   "(what a dense implementation would have to multiply)")
            ach1 = 2.0 * dim * float(n_local) * float(n_global) / kern["rowsum_s"] / 1e12 * kern["live_frac"]
            line["roofline_rowsum"] = dict(bound="tensor", kernel="pair_kernel<%d,MODE_ROWSUM>" % dim, achieved=ach1,
                                           peak=pk["bf16_burst"], unit="TFLOP/s", frac=ach1 / pk["bf16_burst"],
                                           us_per_launch=1e6 * kern["rowsum_s"])
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            q_rows = min(n_global, 4096)
            cpu_reference_step(x_all, cls_all, q_rows, threads)
            tt, pp, reps = 0.0, 0.0, 0
            while tt < 10.0 and reps < 20:
                dt, pairs = cpu_reference_step(x_all, cls_all, q_rows, threads)
                tt += dt; pp += pairs; reps += 1
            line["cpu_baseline"] = dict(value=pp / tt, unit=UNIT, cores=threads, kind="port",
                                        sample="%d query pixels x all %d key pixels fwd+bwd, %d repeats, fp32 torch CPU "
                                               "(oracle port of the reference info_nce composition)" % (q_rows, n_global, reps))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
