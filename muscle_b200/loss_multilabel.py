"""Drop-in for the contrastive members of the reference's src/loss_multilabel.py: same names,
signatures and return conventions; the arithmetic runs in libmuscle_b200.so."""
from __future__ import annotations

import torch

from . import _lib
from ._call import call, ptr, require_cuda_f32


class _ImageLevelContrast(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb, label):
        b, dim = emb.shape
        embc = emb.contiguous()
        labc = label.to(torch.float32).contiguous()
        lib = _lib.load()
        ws_bytes = lib.muscle_imc_workspace_bytes(b, dim)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=emb.device)
        loss = torch.empty((), dtype=torch.float32, device=emb.device)
        n_used = torch.empty((), dtype=torch.int32, device=emb.device)
        demb = torch.empty_like(embc) if emb.requires_grad else None
        call("muscle_imc_fwd_bwd", emb.device, embc.data_ptr(), labc.data_ptr(), b, dim, labc.shape[1], 0.1,
             loss.data_ptr(), n_used.data_ptr(), ptr(demb), ws.data_ptr(), ws_bytes)
        ctx.save_for_backward(demb)
        ctx.mark_non_differentiable(n_used)
        return loss, n_used

    @staticmethod
    def backward(ctx, gloss, _gused):
        demb, = ctx.saved_tensors
        return (None if demb is None else demb * gloss), None


def image_level_contrast(emb, label):
    """src/loss_multilabel.py:36-66 (train_mcl.py:193).  emb [B,ch] fp32 CUDA, label [B,20]
    multi-hot.  Returns a 0-dim tensor with grad to `emb`, or python ``0`` when no image has the
    required positives/negatives — the reference's convention (the caller tests torch.is_tensor);
    deciding that costs the one 4-byte device->host read the reference also performs."""
    require_cuda_f32(emb, "emb")
    if not label.is_cuda:
        raise RuntimeError("muscle_b200: `label` must be a CUDA tensor (there is no CPU path)")
    loss, n_used = _ImageLevelContrast.apply(emb, label)
    if int(n_used.item()) == 0:
        return 0
    return loss


class _PixPro(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fm1s, fm2s, box1, box2, max_px):
        b, c, h, w = fm1s.shape
        f1, f2 = fm1s.contiguous(), fm2s.detach().contiguous()
        loss = torch.empty((), dtype=torch.float32, device=f1.device)
        d1 = torch.empty_like(f1) if fm1s.requires_grad else None
        scratch = torch.empty(1, dtype=torch.float64, device=f1.device)
        call("muscle_pixpro_fwd_bwd", f1.device, f1.data_ptr(), f2.data_ptr(), box1.data_ptr(), box2.data_ptr(), b, c,
             h, w, max_px, loss.data_ptr(), ptr(d1), scratch.data_ptr())
        ctx.save_for_backward(d1)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        d1, = ctx.saved_tensors
        return (None if d1 is None else d1 * gloss), None, None, None, None


def PixPro(fm1s, fm2s, coord1s, coord2s):
    """src/loss_multilabel.py:93-105 (train_mcl.py:209).  fm1s, fm2s [B,21,H,W] fp32 CUDA; coords
    [B,4] = (top, left, h, w) long tensors (CPU, as the DataLoader delivers them, or CUDA).  View 2
    is detached (:103).  Returns the 0-dim loss 1 - mean_b mean_pixels cos."""
    require_cuda_f32(fm1s, "fm1s")
    require_cuda_f32(fm2s, "fm2s")
    b, c, h, w = fm1s.shape
    if fm2s.shape != fm1s.shape:
        raise ValueError("the two views must have the same shape")
    c1 = coord1s.detach().to("cpu").long().reshape(b, 4)
    c2 = coord2s.detach().to("cpu").long().reshape(b, 4)
    if not (torch.equal(c1[:, 2:], c2[:, 2:])):
        raise ValueError("overlap boxes of the two views must have equal sizes")
    for cc in (c1, c2):
        if (cc < 0).any() or (cc[:, 0] + cc[:, 2] > h).any() or (cc[:, 1] + cc[:, 3] > w).any():
            raise ValueError("overlap box outside the %dx%d map" % (h, w))
    max_px = int((c1[:, 2] * c1[:, 3]).max())
    boxes = torch.stack([c1, c2]).to(torch.int32).to(fm1s.device, non_blocking=True)
    return _PixPro.apply(fm1s, fm2s, boxes[0], boxes[1], max_px)
