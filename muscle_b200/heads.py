"""The model-side members of the path: CAM head, image embedding + logits, the trainers' CAM
normalisers (and, in pcm.py, the projection + PCM).  Same arithmetic and argument meaning as the
reference expressions they replace; every step is a C-ABI call (no eager fallback)."""
from __future__ import annotations

import torch

from ._call import call, ptr, require_cuda_f32


class _CamHead(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p7, fc_weight):
        require_cuda_f32(p7, "p7")
        require_cuda_f32(fc_weight, "fc_weight")
        b, ch, h, w = p7.shape
        k = fc_weight.shape[0]
        if fc_weight.shape[1] != ch:
            raise ValueError("fc_weight is %s for %d channels" % (tuple(fc_weight.shape), ch))
        p7c, wc = p7.contiguous(), fc_weight.detach().contiguous()
        cam = torch.empty((b, k, h, w), dtype=torch.float32, device=p7.device)
        call("muscle_cam_head_fwd", p7.device, p7c.data_ptr(), wc.data_ptr(), b, ch, h * w, k, cam.data_ptr())
        ctx.save_for_backward(cam, wc)
        ctx.shape = (b, ch, h, w)
        return cam

    @staticmethod
    def backward(ctx, dcam):
        cam, wc = ctx.saved_tensors
        b, ch, h, w = ctx.shape
        dcam = dcam.contiguous()
        dp7 = torch.empty((b, ch, h, w), dtype=torch.float32, device=cam.device)
        call("muscle_cam_head_bwd", cam.device, dcam.data_ptr(), cam.data_ptr(), wc.data_ptr(), None, b, ch, h * w,
             wc.shape[0], dp7.data_ptr())
        return dp7, None


def cam_head(p7, fc_weight):
    """src/MuSCLe.py:242-247 (== :263-268): relu(sum_ch p7 * fc.weight.data) — a 1x1 conv whose
    weight receives NO gradient (the reference reads `.data`).  p7 [B,ch,h,w] -> [B,K,h,w]."""
    return _CamHead.apply(p7, fc_weight)


class _Gap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p7):
        require_cuda_f32(p7, "p7")
        b, ch, h, w = p7.shape
        emb = torch.empty((b, ch), dtype=torch.float32, device=p7.device)
        call("muscle_gap_fwd", p7.device, p7.contiguous().data_ptr(), b, ch, h * w, emb.data_ptr())
        ctx.shape = (b, ch, h, w)
        return emb

    @staticmethod
    def backward(ctx, demb):
        b, ch, h, w = ctx.shape
        demb = demb.contiguous()
        dp7 = torch.empty((b, ch, h, w), dtype=torch.float32, device=demb.device)
        # the CAM-head backward kernel with no CAM gradient writes exactly demb / (h*w)
        dummy_w = torch.zeros((1, ch), dtype=torch.float32, device=demb.device)
        call("muscle_cam_head_bwd", demb.device, None, None, dummy_w.data_ptr(), demb.data_ptr(), b, ch, h * w, 1,
             dp7.data_ptr())
        return dp7


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb, weight):
        require_cuda_f32(emb, "emb")
        require_cuda_f32(weight, "weight")
        b, ch = emb.shape
        k = weight.shape[0]
        embc, wc = emb.contiguous(), weight.contiguous()
        out = torch.empty((b, k), dtype=torch.float32, device=emb.device)
        call("muscle_linear_fwd", emb.device, embc.data_ptr(), wc.data_ptr(), b, ch, k, out.data_ptr())
        ctx.save_for_backward(embc, wc)
        return out

    @staticmethod
    def backward(ctx, dout):
        embc, wc = ctx.saved_tensors
        b, ch = embc.shape
        k = wc.shape[0]
        demb, dw = torch.empty_like(embc), torch.empty_like(wc)
        call("muscle_linear_bwd", embc.device, dout.contiguous().data_ptr(), embc.data_ptr(), wc.data_ptr(), b, ch, k,
             demb.data_ptr(), dw.data_ptr())
        return demb, dw


def image_embedding_logits(p7, fc_weight):
    """src/MuSCLe.py:182-183,240-241: emb = GAP(p7) [B,ch]; logits = fc(emb) [B,K] (no bias; the
    gradient DOES reach fc.weight here)."""
    emb = _Gap.apply(p7)
    return emb, _Linear.apply(emb, fc_weight)


class _SoftmaxNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cams):
        require_cuda_f32(cams, "cams")
        b, c = cams.shape[:2]
        x = cams.contiguous()
        out = torch.empty_like(x)
        call("muscle_cam_softmaxnorm_fwd", x.device, x.data_ptr(), b, c, x[0, 0].numel(), out.data_ptr())
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, dout):
        out, = ctx.saved_tensors
        b, c = out.shape[:2]
        din = torch.empty_like(out)
        call("muscle_cam_softmaxnorm_bwd", out.device, out.data_ptr(), dout.contiguous().data_ptr(), b, c,
             out[0, 0].numel(), din.data_ptr())
        return din


def cam_softmaxnorm(cams):
    """train_mcl.py:30-36: softmax over the fg channels, bg = 1 - max fg.  [B,C,H,W] -> same."""
    return _SoftmaxNorm.apply(cams)


class _MaxNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cams):
        require_cuda_f32(cams, "cams")
        b, c = cams.shape[:2]
        x = cams.contiguous()
        out = torch.empty_like(x)
        lohi = torch.empty((b * c, 2), dtype=torch.float32, device=x.device)
        argmm = torch.empty((b * c, 2), dtype=torch.int64, device=x.device)
        call("muscle_cam_maxnorm_fwd", x.device, x.data_ptr(), b * c, x[0, 0].numel(), out.data_ptr(), lohi.data_ptr(),
             argmm.data_ptr())
        ctx.save_for_backward(x, lohi, argmm)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, lohi, argmm = ctx.saved_tensors
        b, c = x.shape[:2]
        din = torch.empty_like(x)
        call("muscle_cam_maxnorm_bwd", x.device, x.data_ptr(), dout.contiguous().data_ptr(), lohi.data_ptr(),
             argmm.data_ptr(), b * c, x[0, 0].numel(), din.data_ptr())
        return din


def cam_maxnorm(cams):
    """train_mcl.py:21-28: relu, per-(image,class) min-max with -1e-6 / +1e-6, relu."""
    return _MaxNorm.apply(cams)


class _Pcm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cam, f, fuse_weight, fuse_bias):
        from . import _lib
        lib = _lib.load()
        b, k, h, w = cam.shape
        cin = f.shape[1]
        P = h * w
        camc, fc = cam.contiguous(), f.contiguous()
        wc = fuse_weight.reshape(fuse_weight.shape[0], -1).contiguous()
        bc = fuse_bias.contiguous()
        state_bytes = lib.muscle_pcm_state_bytes(b, cin, P)
        state = torch.empty(state_bytes, dtype=torch.uint8, device=cam.device)
        out = torch.empty((b, k, h, w), dtype=torch.float32, device=cam.device)
        call("muscle_pcm_fwd", cam.device, camc.data_ptr(), fc.data_ptr(), wc.data_ptr(), bc.data_ptr(), b, k, cin, P,
             out.data_ptr(), state.data_ptr(), state_bytes)
        ctx.save_for_backward(camc, out, state)
        ctx.dims = (b, k, cin, h, w)
        ctx.w_shape = fuse_weight.shape
        return out

    @staticmethod
    def backward(ctx, dout):
        from . import _lib
        lib = _lib.load()
        camc, out, state = ctx.saved_tensors
        b, k, cin, h, w = ctx.dims
        P = h * w
        need_cam, _, need_w, need_b = ctx.needs_input_grad
        dev = camc.device
        dcam = torch.empty_like(camc) if need_cam else None
        dw = db = None
        if need_w or need_b:
            dw = torch.empty((128, cin), dtype=torch.float32, device=dev)
            db = torch.empty((128,), dtype=torch.float32, device=dev)
        scratch_bytes = lib.muscle_pcm_bwd_scratch_bytes(b, cin, P)
        scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device=dev)
        call("muscle_pcm_bwd", dev, dout.contiguous().data_ptr(), camc.data_ptr(), out.data_ptr(), b, k, cin, P,
             state.data_ptr(), state.numel(), ptr(dcam), ptr(dw), ptr(db), scratch.data_ptr(), scratch_bytes)
        return dcam, None, (dw.reshape(ctx.w_shape) if need_w else None), (db if need_b else None)


def pcm(cam, f, fuse_weight, fuse_bias):
    """src/MuSCLe.py:213-223 (`MuSCLe.PCM(cam, f)` with the `fuse` 1x1 conv of :181 passed in):
    g = fuse(f); ghat = g/(||g||+1e-5); S = relu(ghat^T ghat); out = cam @ (S / (colsum S + 1e-5)).
    cam [B,K,h,w] (grad), f [B,Cin,h',w'] (treated as detached, :254), fuse_weight [128,Cin,1,1] or
    [128,Cin] and fuse_bias [128] (grad).  When cam and f differ in size the reference first resizes
    cam bilinearly (:215); that resize stays a torch call, the projection, the affinity and the
    aggregation run on the tensor pipe and the P x P affinity never exists in memory."""
    require_cuda_f32(cam, "cam")
    require_cuda_f32(f, "f")
    require_cuda_f32(fuse_weight, "fuse_weight")
    require_cuda_f32(fuse_bias, "fuse_bias")
    if fuse_weight.shape[0] != 128:
        raise ValueError("the projection width is 128 (src/MuSCLe.py:181), got %d" % fuse_weight.shape[0])
    if cam.shape[2:] != f.shape[2:]:
        cam = torch.nn.functional.interpolate(cam, f.shape[2:], mode="bilinear", align_corners=True)
    return _Pcm.apply(cam, f.detach(), fuse_weight, fuse_bias)
