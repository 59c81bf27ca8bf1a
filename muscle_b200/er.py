"""the ER loss of the MCL trainer (train_mcl.py:175-188) and the bilinear up-sampling in front
of it (src/MuSCLe.py:256-257), as autograd ops over libmuscle_b200.so (csrc/er.cu)."""
from __future__ import annotations

import torch

from . import _lib
from ._call import call, require_cuda_f32


class _Upsample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, out_h, out_w):
        xc = x.contiguous()
        lead, (h, w) = xc.shape[:-2], xc.shape[-2:]
        planes = 1
        for d in lead:
            planes *= d
        y = torch.empty(tuple(lead) + (out_h, out_w), dtype=torch.float32, device=xc.device)
        call("muscle_upsample_bilinear_fwd", xc.device, xc.data_ptr(), planes, h, w, out_h, out_w, y.data_ptr())
        ctx.dims = (planes, h, w, out_h, out_w, tuple(xc.shape))
        return y

    @staticmethod
    def backward(ctx, dy):
        planes, h, w, out_h, out_w, shape = ctx.dims
        g = dy.contiguous().float()
        lib = _lib.load()
        sb = int(lib.muscle_upsample_bilinear_bwd_scratch_bytes(planes, out_h, w))
        scratch = torch.empty(sb, dtype=torch.uint8, device=g.device)
        dx = torch.empty(shape, dtype=torch.float32, device=g.device)
        call("muscle_upsample_bilinear_bwd", g.device, g.data_ptr(), planes, h, w, out_h, out_w, dx.data_ptr(),
             scratch.data_ptr(), sb)
        return dx, None, None


def upsample_bilinear(x, size):
    """``F.interpolate(x, size=size, mode='bilinear', align_corners=True)`` (src/MuSCLe.py:256-257, :285) for
    [..., h, w] fp32 CUDA tensors, forward and exact adjoint."""
    require_cuda_f32(x, "x")
    return _Upsample.apply(x, int(size[0]), int(size[1]))


class _ErLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw_cams, raw_sgcs, label):
        cams, sgcs = raw_cams.detach().contiguous(), raw_sgcs.contiguous()
        lab = label.contiguous().float()
        n, c, h, w = sgcs.shape
        lib = _lib.load()
        sb = int(lib.muscle_er_loss_state_bytes(n))
        state = torch.empty(sb, dtype=torch.uint8, device=sgcs.device)
        loss = torch.empty((), dtype=torch.float32, device=sgcs.device)
        call("muscle_er_loss_fwd", sgcs.device, cams.data_ptr(), sgcs.data_ptr(), lab.data_ptr(), n, c, h, w, loss.data_ptr(),
             state.data_ptr(), sb)
        ctx.save_for_backward(cams, sgcs, lab, state)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        cams, sgcs, lab, state = ctx.saved_tensors
        n, c, h, w = sgcs.shape
        g = gloss.detach().reshape(1).contiguous().float()
        dsg = torch.empty_like(sgcs)
        call("muscle_er_loss_bwd", sgcs.device, cams.data_ptr(), sgcs.data_ptr(), lab.data_ptr(), n, c, h, w, state.data_ptr(),
             state.numel(), g.data_ptr(), dsg.data_ptr())
        return None, dsg, None


def er_loss(raw_cams, raw_sgcs, label):
    """train_mcl.py:175-188: ``topk(|cam_softmaxnorm(raw_cams).detach() - cam_softmaxnorm(raw_sgcs)| * label mask,
    k = int(0.2 * int(label.sum()) * h * w)).mean()`` for [B, 21, H, W] maps and [B, 20] multi-hot labels.  Gradient
    reaches ``raw_sgcs`` only, like the reference; nothing is read back to the host."""
    require_cuda_f32(raw_cams, "raw_cams")
    require_cuda_f32(raw_sgcs, "raw_sgcs")
    if raw_cams.shape != raw_sgcs.shape or raw_cams.dim() != 4 or raw_cams.shape[1] != 21:
        raise ValueError("expected raw_cams, raw_sgcs [B, 21, H, W]")
    return _ErLoss.apply(raw_cams, raw_sgcs, label.to(raw_sgcs.device))
