"""Rank program of tests/test_gpu_dist_nccl.py: the batch-sharded dense InfoNCE over NCCL against the
single-GPU op on the concatenated batch and the CPU oracle.  Launched with torch.distributed.run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn.functional as F


def main():
    import muscle_b200 as M
    from muscle_b200 import dist as MD
    from oracle import muscle_oracle as O
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for (gb, d, h, w) in ((2 * world, 128, 16, 16), (world, 64, 20, 12), (2 * world, 256, 16, 16)):
        x, cls, _ = O.synth_embeddings(gb, d, h, w, seed=40 + d)
        lb = gb // world
        xl = x[rank * lb:(rank + 1) * lb].to(dev).requires_grad_()
        cl = cls[rank * lb:(rank + 1) * lb].to(dev)
        loss = MD.dense_infonce_sharded(xl, cl, 0.1, 21)
        loss.backward()
        # single-GPU op on the whole batch (every rank computes it: same device code path)
        xa = x.to(dev).requires_grad_()
        la = M.dense_infonce(xa, cls.to(dev), 0.1, 21)
        la.backward()
        ga = xa.grad[rank * lb:(rank + 1) * lb]
        cos = F.cosine_similarity(xl.grad.flatten().double(), ga.flatten().double(), dim=0).item()
        rel = abs(loss.item() - la.item()) / abs(la.item())
        nrm = abs(xl.grad.norm().item() / ga.norm().item() - 1)
        good = rel < 1e-5 and cos > 0.99999 and nrm < 1e-3
        if rank == 0:
            xr = x.clone().requires_grad_()
            ref = O.dense_infonce_from_features(xr.permute(0, 2, 3, 1).reshape(-1, d), cls.reshape(-1), 0.1, 21)
            gref, = torch.autograd.grad(ref, xr)
            rel_o = abs(loss.item() - ref.item()) / abs(ref.item())
            cos_o = F.cosine_similarity(xl.grad.cpu().flatten().double(), gref[:lb].flatten().double(), dim=0).item()
            good = good and rel_o < 1e-3 and cos_o > 0.999
            print("case gb=%d d=%d: vs 1-GPU rel %.2e cos %.7f | vs oracle rel %.2e cos %.6f -> %s"
                  % (gb, d, rel, cos, rel_o, cos_o, "ok" if good else "FAIL"), flush=True)
        ok = ok and good
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(t) == 1 else 1)


if __name__ == "__main__":
    main()
