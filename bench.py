#!/usr/bin/env python
"""Benchmark of the contrastive-loss hot path (BASELINE.json metric: dense pixel-pair InfoNCE
fwd+bwd, pixel-pairs/s).

    python bench.py --gpus N --steps K --warmup W          # this repo's CUDA path
    python bench.py --impl reference --gpus N ...          # the reference algorithm on host cores

One "step" = one forward + input-gradient pass of the dense label-masked InfoNCE over one batch of
synthetic EfficientNet-stride-8 embeddings.  The workload is THE SAME at every N (so that the driver's
1 -> 8 efficiency is one strong-scaling curve): BASELINE config 4 — global batch 64, 512x512 crop ->
64x64 pixels, D = 128, 21 classes, 262 144 pixels, 6.9e10 pixel pairs — sharded by image over the N
ranks, cross-GPU negatives through one all-gather (N = 1: the single-GPU op on the whole batch).
The N = 1 line also carries config 2 (the configuration the metric is quoted on: B = 8, 448 crop ->
56x56) and config 3 (the largest single-GPU configuration: B = 16, 512 crop -> 64x64) as keyed
sub-records under "configs", each with its own roofline / e2e / cpu_baseline.
`--workload configN` benches one configuration alone.  Prints ONE JSON line on rank 0.

What the line counts: `value` = ALL (query, key) entries of the global similarity matrix per second
(SURVEY.md §8d's pixel pair: N_g^2 per step); pairs of equal class are masked by the loss and cost
nothing on either arm, so `live_pair_fraction` / `live_pairs_per_s` state the cross-class share, and
`class_histogram` the mix it comes from (`config.bg_bias` is the generator's one free parameter).
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

import workloads  # noqa: E402

METRIC = "contrastive_loss_fwd_bwd_pixel_pairs_per_s"
UNIT = "pixel-pairs/s"
TEMPERATURE = 0.1
NUM_CLASSES = 21
SEED = 1234

WORKLOADS = {
    "config2": dict(name="config2: MCL dense InfoNCE fwd+bwd, B=8, 448x448 crop, stride 8 (56x56), D=128, 21 classes",
                    global_batch=8, h=56, w=56, dim=128),
    "config3": dict(name="config3: dense pseudo-mask pixel-pixel InfoNCE fwd+bwd, batch 16, 512x512 crop, stride 8 "
                         "(64x64), D=128, 21 classes", global_batch=16, h=64, w=64, dim=128),
    "config4": dict(name="config4: dense InfoNCE with cross-GPU negatives (all-gather), global batch 64, 512x512 crop, "
                         "stride 8 (64x64), D=128, 21 classes", global_batch=64, h=64, w=64, dim=128),
    "config5": dict(name="config5 (stress): full pixel-pixel dense InfoNCE with cross-GPU negatives, global batch 256, "
                         "512x512 crop, stride 8 (64x64), D=256, 21 classes", global_batch=256, h=64, w=64, dim=256),
}
DEFAULT_WORKLOAD = "config4"      # at EVERY --gpus N
SUB_WORKLOADS = ("config2", "config3")


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

    def __init__(self, index, period_s=0.02):
        # NVML clock / event-reason queries take ~2 ms each and serialise in the driver across processes: polled
        # flat out from every rank they showed up in the 8-rank step time, so ONE rank samples, every 20 ms
        self.index = index
        self.period_s = period_s
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
            self._stop.wait(self.period_s)

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


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm on the host cores (checker code: the only use of oracle/ here)
# ------------------------------------------------------------------------------------------------

def cpu_query_sample(cls, q_rows, top=3):
    """The bounded sample of query pixels one CPU step covers.  The reference composition is one info_nce
    call per class, and every call first gathers its negatives E[cls != c] — a fixed cost the full run
    spreads over ALL queries of the class.  A sample spread thinly over all ~20 classes would pay that
    gather 20 times for a few hundred queries each and understate the CPU arm, so the sample is drawn
    (evenly spread, in proportion) from the `top` largest classes, which hold most of the pixels."""
    c = cls.reshape(-1)
    n = c.numel()
    q_rows = min(n, q_rows)
    hist = torch.bincount(c[c >= 0], minlength=NUM_CLASSES)
    classes = torch.argsort(hist, descending=True)[:top]
    classes = classes[hist[classes] > 0]
    tot = int(hist[classes].sum())
    picks = []
    for ci in classes.tolist():
        idx = torch.nonzero(c == ci).reshape(-1)
        k = max(1, min(idx.numel(), int(round(q_rows * idx.numel() / tot))))
        picks.append(idx[(torch.arange(k, dtype=torch.int64) * idx.numel()) // k])
    return torch.cat(picks)


def cpu_reference_step(x, cls, q_idx, threads):
    """One fwd+bwd of the reference composition for the sampled queries against ALL keys, fp32: one
    reference-style info_nce(E[cls==c], E[cls==c], E[cls!=c]) call per class (oracle port of
    src/loss_multilabel.py:9-21), so — like the GPU arm — only cross-class pairs are
    multiplied.  -> seconds, pixel pairs covered (sampled queries x all keys), loss share of the sample."""
    from oracle import muscle_oracle as O
    torch.set_num_threads(threads)
    d = x.shape[1]
    flat = x.permute(0, 2, 3, 1).reshape(-1, d).clone().requires_grad_()
    c = cls.reshape(-1)
    t0 = time.perf_counter()
    e = O.l2_normalize_pixels(flat)
    loss = O.dense_infonce_per_class(e, c, TEMPERATURE, query_idx=q_idx)
    loss.backward()
    dt = time.perf_counter() - t0
    return dt, float(q_idx.numel()) * flat.shape[0], float(loss.detach())


def cpu_sample_rows(n):
    """Queries per CPU step: ~1e9 pairs (a few seconds on 16 host threads), at most the whole batch."""
    return int(max(256, min(n, (1 << 30) // n)))


def cpu_baseline_record(x, cls, budget_s, max_reps, kind_note):
    threads = os.cpu_count() or 1
    n = x.shape[0] * x.shape[2] * x.shape[3]
    q_idx = cpu_query_sample(cls, cpu_sample_rows(n))
    cpu_reference_step(x, cls, q_idx, threads)                        # warm-up
    tt, pp, reps = 0.0, 0.0, 0
    while tt < budget_s and reps < max_reps:
        dt, pairs, _ = cpu_reference_step(x, cls, q_idx, threads)
        tt += dt
        pp += pairs
        reps += 1
    return dict(value=pp / tt, unit=UNIT, cores=threads, kind="port",
                sample="%d query pixels (evenly spread over the 3 largest classes) x all %d key pixels fwd+bwd, %d repeats, fp32 torch CPU; %s"
                       % (q_idx.numel(), n, reps, kind_note))


CPU_NOTE = ("oracle port of the reference info_nce called once per class (src/loss_multilabel.py:9-21 composed per "
            "SURVEY.md 8c): cross-class pairs only, like the GPU arm; the reference is Python and has no package to "
            "install, and /root/reference does not exist on the GPU box")


def run_reference(args, wl, bg_bias):
    """--impl reference: the reference algorithm timed on this box's host cores; every step is a bounded
    sample of the workload (a few seconds), W warm-up steps then exactly K timed steps."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    x, cls, _ = workloads.synth_embeddings(wl["global_batch"], wl["dim"], wl["h"], wl["w"], SEED, bg_bias=bg_bias)
    n = x.shape[0] * x.shape[2] * x.shape[3]
    q_idx = cpu_query_sample(cls, cpu_sample_rows(n))
    for _ in range(args.warmup):
        cpu_reference_step(x, cls, q_idx, threads)
    tot_t, tot_p = 0.0, 0.0
    for _ in range(args.steps):
        dt, pairs, _ = cpu_reference_step(x, cls, q_idx, threads)
        tot_t += dt
        tot_p += pairs
    v = tot_p / tot_t
    lf = workloads.live_pair_fraction(cls, NUM_CLASSES)
    sample = "%d query pixels (evenly spread over the 3 largest classes) x all %d key pixels per step, fp32 torch CPU; %s" % (q_idx.numel(), n, CPU_NOTE)
    line = dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=1e3 * tot_t / args.steps, higher_is_better=True,
                scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["name"]), bg_bias=bg_bias,
                live_pair_fraction=lf, live_pairs_per_s=v * lf,
                cpu_baseline=dict(value=v, unit=UNIT, cores=threads, kind="port", sample=sample),
                e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

class Env:
    """Process-wide state of the GPU arm."""

    def __init__(self, args):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        # stdout carries the ONE JSON line: everything else a library prints there (NCCL's version banner, say) is
        # sent to stderr for the whole run, and the line is written to the saved descriptor at the end
        sys.stdout.flush()
        self.stdout_fd = os.dup(1)
        os.dup2(2, 1)
        if self.world > 1:
            # NCCL prints its version banner on STDOUT at NCCL_DEBUG=VERSION; stdout carries the one JSON line
            if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=self.dev)   # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed_block(self, fn, steps):
        """K steps back to back between barrier + synchronize on both sides (the bench contract), CUDA events on the
        launching stream; for workloads whose per-step working set exceeds the L2, so no flush is needed. -> seconds"""
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        self.barrier()
        return e0.elapsed_time(e1) / 1e3

    def timed(self, fn, steps):
        """Per-step CUDA-event timing on the launching stream, L2 flushed (untimed) before each step. -> seconds"""
        evs = []
        for _ in range(steps):
            self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / 1e3

    def max_over_ranks(self, values):
        t = torch.tensor(values, dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t]


def traffic_for(key, kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture of this workload."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tpath):
        return None, None
    with open(tpath) as f:
        tj = json.load(f)
    rec = tj.get(key)
    if not rec:
        return None, None
    return rec.get("bytes_per_launch", {}).get(kernel), rec.get("source")


def measure(env, key, wl, steps, warmup, bg_bias, e2e_runs_n, with_fused, cpu_budget_s):
    """Benches one workload on env.world ranks.  -> the record (rank 0) or None."""
    import muscle_b200 as M
    from muscle_b200 import dist as MD
    from muscle_b200.stages import CudaStages
    from muscle_b200 import _lib

    world, rank, dev = env.world, env.rank, env.dev
    gb, h, w, dim = wl["global_batch"], wl["h"], wl["w"], wl["dim"]
    assert gb % world == 0, "global batch must divide over the ranks"
    lb = gb // world
    x_all, cls_all, _ = workloads.synth_embeddings(gb, dim, h, w, SEED, bg_bias=bg_bias)
    x_h = x_all[rank * lb:(rank + 1) * lb].contiguous().pin_memory()
    c_h = cls_all[rank * lb:(rank + 1) * lb].contiguous().pin_memory()
    x_d = x_h.to(dev)
    c_d = c_h.to(dev)
    n_local = lb * h * w
    n_global = gb * h * w
    pairs_per_step = float(n_global) * float(n_global)     # every (query, key) entry of the global matrix
    live_frac_pairs = workloads.live_pair_fraction(cls_all, NUM_CLASSES)

    def loss_of(xg, cg):
        if world > 1:
            return MD.dense_infonce_sharded(xg, cg, TEMPERATURE, NUM_CLASSES)
        return M.dense_infonce(xg, cg, TEMPERATURE, NUM_CLASSES)

    last = {}

    def step_device():
        xg = x_d.detach().requires_grad_()
        loss = loss_of(xg, c_d)
        loss.backward()
        last["loss"] = loss.detach()

    def step_e2e():
        xg = x_h.to(dev, non_blocking=True).requires_grad_()
        cg = c_h.to(dev, non_blocking=True)
        loss = loss_of(xg, cg)
        loss.backward()
        return float(loss.item())          # device -> host read of the step's result

    # bytes one step touches on this rank: fp32 input + gradient, gathered bf16 keys, the gradient pass's partial slots
    part_bytes = 8 * ((n_local + 255) // 256 * 256) * dim * 4
    working_set = 2 * n_local * dim * 4 + n_global * dim * 2 + part_bytes
    big = working_set > 1.5 * 126 * 1024 * 1024
    # warm-up: the W steps asked for, and at least 10 at N > 1 (NCCL sets its channels up lazily and the SM clock
    # needs a few tens of milliseconds of load to settle; the same count on every rank: the steps hold collectives)
    warm_run = max(warmup, 10) if world > 1 else warmup
    for _ in range(warm_run):
        step_device()
    env.barrier()
    # rank 0's GPU speaks for the box
    sampler = ClockSampler(env.local_rank, 0.02 if world > 1 else 0.004) if rank == 0 else None
    if sampler is not None:
        sampler.start()
    gc_was = gc.isenabled()
    gc.disable()                      # (as in the end-to-end loop below: a collection costs several short steps)
    if big:
        # three blocks of EXACTLY K steps, each bracketed by barrier + synchronize; the line reports the median block
        # (max over ranks per block) and lists all three: one 8-rank block is ~0.1 s, and a single host or NCCL
        # hiccup on any rank moved it by 10 % from run to run
        blocks = env.max_over_ranks([env.timed_block(step_device, steps) for _ in range(3)])
        t_dev = sorted(blocks)[1]
    else:
        blocks = None
        t_dev = env.timed(step_device, steps)
    if gc_was:
        gc.enable()
    env.barrier()
    clocks = sampler.stop() if sampler is not None else None
    loss_value = float(last["loss"].item())
    for _ in range(3):
        step_e2e()
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step_e2e()
    torch.cuda.synchronize()
    t_e2e_serial = time.perf_counter() - t0
    env.barrier()

    # The same through the public API the way a training loop feeds it: every step's inputs still go
    # pinned host -> device and every step's loss still comes back to the host inside the timed region,
    # but the next step's upload runs on a copy stream under the current step's kernels and the loss is
    # read back asynchronously (checked one step later), so no step waits on PCIe or on a host sync.
    copy_stream = torch.cuda.Stream(device=dev)
    loss_pinned = torch.empty(steps + 3, dtype=torch.float32).pin_memory()
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

    def e2e_pipelined(n_steps, fused=False):
        main = torch.cuda.current_stream()
        for e in use_ev:
            e.record(main)
        gc_was = gc.isenabled()
        gc.disable()                  # a collection in the middle of a short step costs several steps
        t_host0 = time.perf_counter()
        upload(0)
        for i in range(n_steps):
            b = i & 1
            main.wait_event(up_ev[b])
            if i + 1 < n_steps:
                upload(i + 1)
            if fused:       # loss and dx from one C call, no autograd (dense_infonce_value_and_grad)
                loss, _dx = M.dense_infonce_value_and_grad(x_buf[b], c_buf[b], TEMPERATURE, NUM_CLASSES)
            else:
                xg = x_buf[b].detach().requires_grad_()
                loss = loss_of(xg, c_buf[b])
                loss.backward()
            use_ev[b].record(main)
            loss_pinned[i].copy_(loss.detach(), non_blocking=True)   # device -> host read of the step's result
        e2e_pipelined.host_s = time.perf_counter() - t_host0          # all steps enqueued (host side only)
        torch.cuda.synchronize()
        if gc_was:
            gc.enable()
        return loss_pinned[:n_steps].tolist()

    e2e_pipelined(3)
    e2e_runs = []
    for _ in range(e2e_runs_n):
        env.barrier()
        t0 = time.perf_counter()
        e2e_losses = e2e_pipelined(steps)
        e2e_runs.append((time.perf_counter() - t0, e2e_pipelined.host_s))
        assert all(v == v for v in e2e_losses), "NaN loss in the end-to-end run"
    e2e_runs.sort()
    t_e2e, e2e_host_s = e2e_runs[len(e2e_runs) // 2]
    env.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        upload(i)
    torch.cuda.synchronize()
    t_h2d = time.perf_counter() - t0
    fused_e2e = None
    if world == 1 and with_fused:      # informational: the same loop through the autograd-free public call
        e2e_pipelined(3, fused=True)
        fr = []
        for _ in range(e2e_runs_n):
            env.barrier()
            t0 = time.perf_counter()
            e2e_pipelined(steps, fused=True)
            fr.append((time.perf_counter() - t0, e2e_pipelined.host_s))
        fr.sort()
        mid = fr[len(fr) // 2]
        fused_e2e = dict(ms_per_step=1e3 * mid[0] / steps, value=pairs_per_step * steps / mid[0],
                         host_enqueue_ms_per_step=1e3 * mid[1] / steps,
                         note="dense_infonce_value_and_grad: loss and dx from one C call, no autograd engine")
    env.barrier()
    del x_buf, c_buf

    # ---- the two pair passes timed alone through the staged C ABI (this rank's rows x all columns)
    st = CudaStages(dev)
    C = NUM_CLASSES
    f32, i32 = torch.float32, torch.int32
    xg = x_d.detach().requires_grad_()
    loss0, state = MD.sharded_fwd(st, xg, c_d, TEMPERATURE, C, 1e-12, True, None)
    rows_pad, cols_pad = state["rows_pad"], state["cols_pad"]
    part = st.empty((st.part_slots(), rows_pad, dim), f32)
    w_cols = state["w_all"] if world > 1 else state["w"]

    def grad_fn():
        st.grad(state["ehat"], state["bits"], state["tor"], state["w"], rows_pad, state["ehat_all"], state["bits_all"],
                state["tor_all"], w_cols, cols_pad, state["unit_off"], dim, TEMPERATURE, part)

    rs_tmp = st.zeros((rows_pad,), f32)
    sym = dim <= 128 and os.environ.get("MUSCLE_ROWSUM_SYM", "1") != "0"
    if sym:
        rs_all = st.zeros((cols_pad,), f32)
        if world > 1 and os.environ.get("MUSCLE_ROWSUM_RING", "0") != "1":    # as muscle_b200.dist does it: split by count
            uo_sym = st.empty((cols_pad // 128 + 1,), i32)
            st.plan_sym_ring(state["tor_all"], cols_pad, state["tor_all"], cols_pad, 0, 1, uo_sym)

            def rowsum_fn():
                st.rowsum_sym_share(state["ehat_all"], state["bits_all"], state["tor_all"], cols_pad, uo_sym, rank, world,
                                    dim, TEMPERATURE, rs_all)
        else:
            uo_sym = st.empty((rows_pad // 128 + 1,), i32)
            st.plan_sym_ring(state["tor"], rows_pad, state["tor_all"], cols_pad, rank, world, uo_sym)

            def rowsum_fn():
                st.rowsum_sym_ring(state["ehat"], state["bits"], state["tor"], rows_pad, state["ehat_all"],
                                   state["bits_all"], state["tor_all"], cols_pad, uo_sym, rank, world, dim, TEMPERATURE,
                                   rs_all)
        rowsum_kernel = "pair_sym_kernel<%d>" % dim
    else:
        def rowsum_fn():
            st.rowsum(state["ehat"], state["bits"], state["tor"], rows_pad, state["ehat_all"], state["bits_all"],
                      state["tor_all"], cols_pad, state["unit_off"], dim, TEMPERATURE, rs_tmp)
        rowsum_kernel = "pair_kernel<%d,MODE_ROWSUM>" % dim
    live_units = int(state["unit_off"][-1].item())
    for _ in range(3):
        grad_fn()
        rowsum_fn()
    ksteps = max(5, min(steps, 30))
    grad_s = env.timed(grad_fn, ksteps) / ksteps
    rowsum_s = env.timed(rowsum_fn, ksteps) / ksteps
    live_tile_frac = live_units / float((rows_pad // 128) * (cols_pad // 128))
    del part, state

    t_dev, t_e2e, t_e2e_serial, grad_s, rowsum_s = env.max_over_ranks([t_dev, t_e2e, t_e2e_serial, grad_s, rowsum_s])
    live_units_all = env.max_over_ranks([float(live_units)])[0]     # slowest rank's share (the one timed)
    if rank != 0:
        return None

    pk = peaks()
    value = pairs_per_step * steps / t_dev
    rec = dict(workload=wl["name"], value=value, unit=UNIT, ms_per_step=1e3 * t_dev / steps, steps=steps, warmup=warmup,
               warmup_steps_run=warm_run, timed_blocks_ms_per_step=(None if blocks is None else [1e3 * b / steps for b in blocks]),
               loss=loss_value, live_pair_fraction=live_frac_pairs, live_pairs_per_s=value * live_frac_pairs,
               class_histogram=workloads.class_histogram(cls_all, NUM_CLASSES),
               global_batch=gb, pixels=n_global, pixels_per_image=h * w, proj_dim=dim, clocks=clocks,
               l2=("a step touches %.0f MB on every rank (input, gradient, gathered keys, partial-gradient slots): well "
                   "above the 126 MB L2, K steps timed back to back between barriers, no flush" % (working_set / 1e6)
                   if big else "flushed (256 MiB write) before every timed step, per-step CUDA events"))
    h2d_bytes = int(x_h.numel() * 4 + c_h.numel() * 8)
    rec["e2e"] = dict(value=pairs_per_step * steps / t_e2e, unit=UNIT, h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=4,
                      ms_per_step=1e3 * t_e2e / steps,
                      mode="pipelined: next step's pinned-host upload on a copy stream into one of two device buffers, "
                           "loss read back asynchronously; every step's H2D and D2H are inside the timed region",
                      host_enqueue_ms_per_step=1e3 * e2e_host_s / steps, value_and_grad=fused_e2e,
                      h2d_only_ms_per_step=1e3 * t_h2d / steps, h2d_gb_per_s=h2d_bytes * steps / t_h2d / 1e9,
                      runs_ms_per_step=[1e3 * r[0] / steps for r in e2e_runs],
                      runs_note="%d run(s) of `steps` steps each, the median is `value`" % len(e2e_runs),
                      serial_value=pairs_per_step * steps / t_e2e_serial, serial_ms_per_step=1e3 * t_e2e_serial / steps,
                      serial_note="upload -> step -> blocking .item() per step, nothing overlapped")
    # whole step on the pairs the loss keeps: 6*D flops per ordered LIVE pixel pair (S, dQ, dK) / step time
    live_pairs = pairs_per_step * live_frac_pairs
    live_tf = 6.0 * dim * live_pairs * steps / t_dev / 1e12 / world
    all_tf = 6.0 * dim * pairs_per_step * steps / t_dev / 1e12 / world
    rec["step_tensor"] = dict(live_tflops_per_gpu=live_tf, live_frac_of_sustained=live_tf / pk["bf16_sustained"],
                              live_frac_of_burst=live_tf / pk["bf16_burst"], all_pairs_tflops_per_gpu=all_tf,
                              note="per GPU, whole step incl. HBM kernels and collectives; 6*D flops per ordered pixel "
                                   "pair of different class (the pairs the loss keeps); all_pairs counts the masked "
                                   "same-class pairs too, which no arm multiplies")
    # dominant kernel: the gradient pass.  Algorithmic flops of a launch = 4*D per ordered pair of every LIVE
    # 128 x 128 tile pair (tile pairs of one single class are masked entirely and never multiplied)
    unit_pairs = live_units_all * 128.0 * 128.0
    ach = 4.0 * dim * unit_pairs / grad_s / 1e12
    traffic, traffic_src = traffic_for(key, "pair_kernel<%d,MODE_GRAD>" % dim) if world == 1 else (None, None)
    rec["roofline"] = dict(bound="tensor", kernel="pair_kernel<%d,MODE_GRAD>" % dim, achieved=ach, peak=pk["bf16_burst"],
                           unit="TFLOP/s", frac=ach / pk["bf16_burst"], frac_of_sustained=ach / pk["bf16_sustained"],
                           traffic=traffic, traffic_source=traffic_src,
                           traffic_note="algorithmic DRAM bytes per launch = the bf16 operands once: %d bytes"
                                        % ((rows_pad + (cols_pad if world > 1 else 0)) * dim * 2),
                           peak_source=pk["source"] + ", burst figure (kernel timed alone)",
                           algorithmic="4*D flops per ordered pixel pair of every live (row tile, column tile) unit: S "
                                       "recomputed (2*D) + one T@E product (2*D) that serves dQ and dK through the "
                                       "symmetric form; slowest rank's launch",
                           us_per_launch=1e6 * grad_s, live_tile_fraction=live_tile_frac, live_units=int(live_units_all))
    ach1 = 2.0 * dim * unit_pairs / rowsum_s / 1e12
    traffic1, traffic1_src = traffic_for(key, rowsum_kernel) if world == 1 else (None, None)
    rec["roofline_rowsum"] = dict(bound="tensor", kernel=rowsum_kernel, achieved=ach1, peak=pk["bf16_burst"], unit="TFLOP/s",
                                  frac=ach1 / pk["bf16_burst"], us_per_launch=1e6 * rowsum_s, traffic=traffic1,
                                  traffic_source=traffic1_src,
                                  algorithmic="2*D flops per ordered pixel pair of every live unit (S = E E^T); the "
                                              "symmetric pass executes half of them (S of a tile pair serves both "
                                              "orders), so frac can exceed the executed tensor rate" if sym else
                                              "2*D flops per ordered pixel pair of every live unit (S = R C^T)")
    if cpu_budget_s > 0:
        rec["cpu_baseline"] = cpu_baseline_record(x_all, cls_all, cpu_budget_s, 20, CPU_NOTE)
    return rec


def pool_record(env, n, dim, num_classes, steps):
    """HBM roofline of the CAM-weighted prototype pooling kernels (csrc/pool.cu) at one workload's size: bf16
    class-sorted embeddings [n, dim] + fp32 weights [n, C]; algorithmic bytes per SURVEY.md 8d."""
    import muscle_b200 as M
    from muscle_b200 import _lib
    from muscle_b200._call import call
    dev = env.dev
    gen = torch.Generator(device=dev).manual_seed(SEED)
    e = torch.nn.functional.normalize(torch.randn(n, dim, device=dev, generator=gen), dim=1).to(torch.bfloat16)
    w = torch.relu(torch.randn(n, num_classes, device=dev, generator=gen))
    sums = torch.empty(num_classes, dim, device=dev)
    mass = torch.empty(num_classes, device=dev)
    proto = torch.empty(num_classes, dim, device=dev)
    dproto = torch.randn(num_classes, dim, device=dev, generator=gen)
    de = torch.empty(n, dim, device=dev)
    dw = torch.empty(n, num_classes, device=dev)
    _lib.load()

    sb = int(_lib.load().muscle_pool_prototypes_fwd_scratch_bytes(dim, num_classes))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)

    def fwd():
        call("muscle_pool_prototypes_fwd_ws", dev, e.data_ptr(), 1, w.data_ptr(), n, dim, num_classes, 1e-30,
             sums.data_ptr(), mass.data_ptr(), proto.data_ptr(), scratch.data_ptr(), sb)

    def bwd():
        call("muscle_pool_prototypes_bwd", dev, e.data_ptr(), 1, w.data_ptr(), n, dim, num_classes, 1e-30, proto.data_ptr(),
             mass.data_ptr(), dproto.data_ptr(), de.data_ptr(), dw.data_ptr())

    for _ in range(3):
        fwd()
        bwd()
    k = max(10, min(steps, 50))
    t_f = env.timed(fwd, k) / k
    t_b = env.timed(bwd, k) / k
    pk = peaks()
    b_f = n * dim * 2 + n * num_classes * 4 + num_classes * dim * 4
    b_b = n * dim * (2 + 4) + n * num_classes * (4 + 4) + 2 * num_classes * dim * 4
    rec = {}
    tkey = "pool_%dx%dx%d" % (n, dim, num_classes)
    for name, t, nbytes, note in (("fwd", t_f, b_f, "N*D*2 (bf16 rows) + N*C*4 (weights) + C*D*4 (SURVEY 8d)"),
                                  ("bwd", t_b, b_b, "N*D*(2 read + 4 write) + N*C*(4 read + 4 write) + 2*C*D*4")):
        ach = nbytes / t / 1e9
        kern = ("pool_tc_fwd_kernel<%d> + pool_tc_finish_kernel" % dim) if name == "fwd" else ("pool_tc_bwd_kernel<%d>" % dim)
        rec[name] = dict(bound="hbm", kernel=kern, achieved=ach, peak=pk["hbm"], unit="GB/s",
                         frac=ach / pk["hbm"], us_per_launch=1e6 * t, algorithmic_bytes=nbytes, algorithmic=note,
                         traffic=traffic_for(tkey, "pool_" + name)[0], peak_source=pk["source"])
    rec["shape"] = dict(pixels=n, dim=dim, num_classes=num_classes, e="bf16", w="f32")
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="N=1: skip the config2 / config3 sub-records")
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="bench this configuration alone (default: config4 at every N, + config2/config3 sub-records at N=1)")
    ap.add_argument("--bg-bias", type=float, default=0.0,
                    help="bias of the background channel in the synthetic class maps (0 = SURVEY 8d; 0.15 = round 1, ~75%% bg)")
    args = ap.parse_args()
    key = args.workload or DEFAULT_WORKLOAD
    wl = WORKLOADS[key]
    if args.impl == "reference":
        return run_reference(args, wl, args.bg_bias)
    args.warmup = max(args.warmup, 3)

    env = Env(args)
    cpu_budget = 0 if (args.no_cpu_baseline or env.world > 1) else 12.0
    rec = measure(env, key, wl, args.steps, args.warmup, args.bg_bias, 3, True, cpu_budget)
    subs = {}
    if env.world == 1 and args.workload is None and not args.no_sub:
        for sk in SUB_WORKLOADS:
            sub_steps = max(args.steps, 20)
            subs[sk] = measure(env, sk, WORKLOADS[sk], sub_steps, args.warmup, args.bg_bias, 3, True,
                               0 if args.no_cpu_baseline else 6.0)
    pool = None
    if env.world == 1 and args.workload is None and not args.no_sub:
        c3 = WORKLOADS["config3"]       # the largest single-GPU configuration: 65 536 pixels, D = 128, 21 classes
        pool = pool_record(env, c3["global_batch"] * c3["h"] * c3["w"], c3["dim"], NUM_CLASSES, args.steps)
        # ... and at the headline workload's size (config 4: 262 144 pixels): at 65 536 pixels the whole input is 22 MB, a
        # few microseconds of HBM time, and launch + pipeline fill dominate the launch
        pool["at_config4_size"] = pool_record(env, wl["global_batch"] * wl["h"] * wl["w"], wl["dim"], NUM_CLASSES, args.steps)
    ops = None
    if env.world == 1 and args.workload is None and not args.no_sub:
        # the other members of the path (PCM, CAM head, IMC, PixPro, EMD, FieldLoss, ER loss, ...): device time fwd+bwd
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_ops
        ops = bench_ops.collect(iters=10, with_cpu=False, big=True)
    if env.rank == 0:
        world = env.world
        launches_per_step = 7 if world == 1 else 11      # ours; NCCL's kernels are not counted
        line = dict(metric=METRIC, value=rec["value"], unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=rec["ms_per_step"], higher_is_better=True, scaling="strong", vs_baseline=None,
                    dtype="bf16", data="synthetic",
                    config=dict(workload=wl["name"], global_batch=wl["global_batch"], pixels_per_image=wl["h"] * wl["w"],
                                proj_dim=wl["dim"], temperature=TEMPERATURE, bg_bias=args.bg_bias,
                                parallelism="dp%d (batch-sharded, all-gather keys)" % world,
                                l2=rec["l2"],
                                accumulate="fp32 (TMEM); operands bf16",
                                same_workload_at_every_n=args.workload is None),
                    loss=rec["loss"], live_pair_fraction=rec["live_pair_fraction"],
                    live_pairs_per_s=rec["live_pairs_per_s"], class_histogram=rec["class_histogram"],
                    e2e=rec["e2e"], gpu_launches=int(args.steps * launches_per_step),
                    gpu_launches_note="%d kernels of libmuscle_b200.so per step (hist, scan, prepare (+ class sums), row-sum "
                                      "pass, mid, gradient pass, final%s)" % (launches_per_step,
                                                                                "" if world == 1 else ", plan, symmetric plan x2, loss"),
                    clocks=rec["clocks"], step_tensor=rec["step_tensor"], roofline=rec["roofline"],
                    roofline_rowsum=rec["roofline_rowsum"])
        if rec.get("timed_blocks_ms_per_step"):
            line["timed_blocks_ms_per_step"] = rec["timed_blocks_ms_per_step"]
            line["timing"] = "median of 3 blocks of K steps, each between barrier + synchronize, max over ranks per block"
        if "cpu_baseline" in rec:
            line["cpu_baseline"] = rec["cpu_baseline"]
        if subs:
            line["configs"] = {k: v for k, v in subs.items() if v is not None}
        if pool is not None:
            line["roofline_pool"] = pool
        if ops is not None:
            line["ops"] = ops
        os.write(env.stdout_fd, (json.dumps(line) + "\n").encode())
    if env.world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
