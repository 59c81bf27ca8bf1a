"""Host-side orchestration of the batch-sharded dense InfoNCE (muscle_b200/dist.py) on CPU:
world_size 1 in-process and world_size 2 over gloo, with the CUDA stages replaced by the CPU
stand-in of tests/cpu_stages.py.  Contract: loss and d(loss)/d(x_local) at G ranks
== the one-rank result on the concatenated batch == the oracle on the concatenated batch."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import muscle_oracle as O  # noqa: E402


def _oracle(x, cls, temperature=0.1, num_classes=21):
    d = x.shape[1]
    xr = x.clone().requires_grad_()
    ref = O.dense_infonce_from_features(xr.permute(0, 2, 3, 1).reshape(-1, d), cls.reshape(-1), temperature,
                                        num_classes)
    g, = torch.autograd.grad(ref, xr)
    return ref.item(), g


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def test_single_rank_stand_in_matches_oracle():
    from cpu_stages import CpuStages
    from muscle_b200 import dist as MD
    x, cls, _ = O.synth_embeddings(2, 64, 12, 14, seed=3)
    cls = cls.clone()
    cls[0, 0, :5] = -1
    loss, dx = MD.sharded_fwd_bwd(CpuStages(), x, cls, 0.1, 21, 1e-12, True)
    ref, gref = _oracle(x, cls)
    assert abs(loss.item() - ref) <= 1e-3 * abs(ref)
    assert _cos(dx, gref) >= 0.999
    assert float(dx[0, :, 0, :5].abs().max()) == 0.0
    loss2, dx2 = MD.sharded_fwd_bwd(CpuStages(), x, cls, 0.1, 21, 1e-12, False)
    assert dx2 is None and abs(loss2.item() - loss.item()) < 1e-6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from cpu_stages import CpuStages
    from muscle_b200 import dist as MD
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x, cls, _ = O.synth_embeddings(4, 64, 8, 16, seed=5)     # 4 images, 2 per rank
        lb = 4 // world
        xl = x[rank * lb:(rank + 1) * lb].contiguous()
        cl = cls[rank * lb:(rank + 1) * lb].contiguous()
        loss, dx = MD.sharded_fwd_bwd(CpuStages(), xl, cl, 0.1, 21, 1e-12, True)
        # pixel-major layout on the same shard must give the same numbers
        xp = xl.permute(0, 2, 3, 1).reshape(-1, 64).contiguous()
        loss_p, dx_p = MD.sharded_fwd_bwd(CpuStages(), xp, cl.reshape(-1), 0.1, 21, 1e-12, True)
        out[rank] = (loss.item(), dx.clone(), loss_p.item(),
                     dx_p.reshape(lb, 8, 16, 64).permute(0, 3, 1, 2).clone())
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world_size_2_gloo_matches_concatenated_batch():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0, "rank exited with %s" % p.exitcode
    x, cls, _ = O.synth_embeddings(4, 64, 8, 16, seed=5)
    ref, gref = _oracle(x, cls)
    lb = 2
    for r in range(world):
        loss, dx, loss_p, dx_p = out[r]
        assert abs(loss - ref) <= 1e-3 * abs(ref), (r, loss, ref)          # the GLOBAL mean on every rank
        assert _cos(dx, gref[r * lb:(r + 1) * lb]) >= 0.999
        assert abs(dx.norm().item() / gref[r * lb:(r + 1) * lb].norm().item() - 1) < 2e-2
        assert abs(loss_p - loss) <= 1e-6 * abs(loss)
        assert _cos(dx_p, dx) >= 0.999999
    assert abs(out[0][0] - out[1][0]) <= 1e-6 * abs(ref)
