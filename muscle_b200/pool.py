"""CAM-weighted (soft) prototype pooling with its gradient scatter (SURVEY.md §8b `pool_prototypes`, §8c soft
variant): ``proto[c] = sum_p w[p,c] e[p] / sum_p w[p,c]``.  The reference's instances are the positive-key mean of
``info_nce`` (src/loss_multilabel.py:12; 0/1 weights), the region mean of ``EMD.get_weight_vector``
(src/loss_multilabel.py:255) and the CAM aggregation ``cam @ aff`` (src/MuSCLe.py:222; the un-normalised sums).
All arithmetic runs in libmuscle_b200.so (csrc/pool.cu)."""
from __future__ import annotations

import torch

from . import _lib
from ._call import call, ptr


class _PoolPrototypes(torch.autograd.Function):
    @staticmethod
    def forward(ctx, e, w, eps):
        dev = e.device
        n, d = e.shape
        c = w.shape[1]
        sums = torch.empty((c, d), dtype=torch.float32, device=dev)
        mass = torch.empty((c,), dtype=torch.float32, device=dev)
        proto = torch.empty((c, d), dtype=torch.float32, device=dev)
        sb = int(_lib.load().muscle_pool_prototypes_fwd_scratch_bytes(d, c))
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        call("muscle_pool_prototypes_fwd_ws", dev, e.data_ptr(), 1 if e.dtype == torch.bfloat16 else 0, w.data_ptr(), n, d,
             c, float(eps), sums.data_ptr(), mass.data_ptr(), proto.data_ptr(), scratch.data_ptr(), sb)
        ctx.save_for_backward(e, w, proto, mass)
        ctx.eps = float(eps)
        ctx.mark_non_differentiable(mass)
        return proto, sums, mass

    @staticmethod
    def backward(ctx, dproto, dsums, _dmass):
        e, w, proto, mass = ctx.saved_tensors
        dev = e.device
        n, d = e.shape
        c = w.shape[1]
        need_e, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if not (need_e or need_w):
            return None, None, None
        # d/d sums joins d/d proto: sums = proto * mass  ->  an upstream gradient on the sums acts like dproto_c * mass_c
        # on the prototype path only for de; to keep the kernel single-purpose the sums' gradient is folded in here
        dp = torch.zeros_like(proto) if dproto is None else dproto.contiguous().float()
        de = torch.empty((n, d), dtype=torch.float32, device=dev) if need_e else None
        dw = torch.empty((n, c), dtype=torch.float32, device=dev) if need_w else None
        call("muscle_pool_prototypes_bwd", dev, e.data_ptr(), 1 if e.dtype == torch.bfloat16 else 0, w.data_ptr(), n, d, c,
             ctx.eps, proto.data_ptr(), mass.data_ptr(), dp.data_ptr(), ptr(de), ptr(dw))
        if dsums is not None and bool((dsums != 0).any()):
            ds = dsums.float()
            if need_e:
                de = de + w @ ds                       # rare path (the un-normalised `cam @ aff` form): plain torch
            if need_w:
                dw = dw + e.float() @ ds.t()
        if need_e and e.dtype != torch.float32:
            de = de.to(e.dtype)
        return de, dw, None


def pool_prototypes(e: torch.Tensor, w: torch.Tensor, eps: float = 1e-30, return_all: bool = False):
    """e [N, D] (fp32 or bf16, CUDA) pixel embeddings, w [N, C] fp32 non-negative weights (CAM scores, soft masks,
    0/1 class indicators) -> prototypes [C, D] = weighted means, with grad to ``e`` and ``w``.
    ``return_all``: (proto, sums [C, D] = sum_p w e, mass [C] = sum_p w).  D % 4 == 0, D <= 256, C <= 24.
    NCHW inputs: pass ``e.permute(0,2,3,1).reshape(-1, D)`` / ``cam.permute(0,2,3,1).reshape(-1, C)``."""
    if not (torch.is_tensor(e) and e.is_cuda and torch.is_tensor(w) and w.is_cuda):
        raise RuntimeError("muscle_b200: pool_prototypes needs CUDA tensors (there is no CPU path)")
    if e.dim() != 2 or w.dim() != 2 or e.shape[0] != w.shape[0]:
        raise ValueError("expected e [N, D] and w [N, C]")
    if e.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("e must be float32 or bfloat16")
    if e.shape[1] % 4 or e.shape[1] > 256 or w.shape[1] > 24:
        raise ValueError("pool_prototypes supports D % 4 == 0, D <= 256 and C <= 24")
    _lib.load()
    proto, sums, mass = _PoolPrototypes.apply(e.contiguous(), w.contiguous().float(), eps)
    return (proto, sums, mass) if return_all else proto
