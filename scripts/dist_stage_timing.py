"""Per-stage device time of the SHARDED dense InfoNCE step (config 4, `bench.py --gpus N`'s workload).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 scripts/dist_stage_timing.py [--steps 20]

CUDA events are recorded on the current stream around every stage call and every collective of
muscle_b200.dist (debugging aid: the event pairs serialise nothing but add ~2 us each).  Rank 0
prints, per stage, the mean over steps of the MAX over ranks, and the step total next to it."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import bench
    import muscle_b200.dist as MD
    from muscle_b200.stages import CudaStages

    import workloads
    wl = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
    per = wl["global_batch"] // world
    x, cls, _ = workloads.synth_embeddings(wl["global_batch"], wl["dim"], wl["h"], wl["w"], bench.SEED, bg_bias=0.0)
    xd = x[rank * per:(rank + 1) * per].to(dev)
    cg = cls[rank * per:(rank + 1) * per].to(dev)

    records = []     # (name, start event, end event) of the current step

    def wrap(owner, name, label):
        fn = getattr(owner, name)

        def timed(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            records.append((label, e0, e1))
            return out
        setattr(owner, name, timed)

    for m in ("prepare", "plan", "plan_sym_ring", "rowsum", "rowsum_sym_ring", "rowsum_sym_share", "mid", "grad", "final"):
        wrap(CudaStages, m, m)
    wrap(MD.dist, "all_reduce", "nccl all_reduce")
    wrap(MD, "_all_gather_into", "nccl all_gather")
    wrap(MD, "_all_gather_many", "nccl all_gather")
    wrap(CudaStages, "zeros", "torch.zeros")

    totals, step_ms = {}, []
    for it in range(args.warmup + args.steps):
        records.clear()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        xg = xd.detach().requires_grad_()
        s0.record()
        loss = MD.dense_infonce_sharded(xg, cg, bench.TEMPERATURE, bench.NUM_CLASSES)
        loss.backward()
        s1.record()
        torch.cuda.synchronize()
        if it < args.warmup:
            continue
        # gaps between consecutive recorded calls (allocations, torch glue, waiting for the host)
        gap = 0.0
        prev_end = s0
        for label, e0, e1 in records:
            gap += max(0.0, prev_end.elapsed_time(e0))
            prev_end = e1
        gap += max(0.0, prev_end.elapsed_time(s1))
        totals.setdefault("gaps between calls", []).append(gap)
        seen = {}
        for label, e0, e1 in records:
            k = seen.get(label, 0)
            seen[label] = k + 1
            key = "%s #%d" % (label, k) if label.startswith("nccl") or label in ("final", "torch.zeros") else label
            totals.setdefault(key, []).append(e0.elapsed_time(e1))
        step_ms.append(s0.elapsed_time(s1))
    keys = list(totals)
    mine = torch.tensor([sum(totals[k]) / len(totals[k]) for k in keys] + [sum(step_ms) / len(step_ms)],
                        dtype=torch.float64, device=dev)
    mx, mn = mine.clone(), mine.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    if rank == 0:
        out = {"world": world, "workload": wl["name"], "steps": args.steps,
               "stages_ms": {k: {"max_over_ranks": round(float(a), 4), "min_over_ranks": round(float(b), 4)}
                             for k, a, b in zip(keys, mx[:-1], mn[:-1])},
               "step_ms": {"max_over_ranks": round(float(mx[-1]), 4), "min_over_ranks": round(float(mn[-1]), 4)},
               "note": "order of the keys is the order of the calls in one step; a collective's time includes waiting "
                       "for the slowest rank to arrive"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
