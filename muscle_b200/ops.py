"""PyTorch-facing ops of the hot path.  PyTorch supplies device memory, streams and autograd
plumbing only; all arithmetic runs in libmuscle_b200.so through raw pointers."""
from __future__ import annotations

import torch

from . import _lib


def _stream_ptr(device):
    return torch.cuda.current_stream(device).cuda_stream


class _on_device:
    """`with torch.cuda.device(d)` only when `d` is not already current (the context manager costs
    ~10 us of host time per call, which matters for a 0.3 ms step)."""

    def __init__(self, device):
        idx = device.index if device.index is not None else torch.cuda.current_device()
        self.ctx = None if idx == torch.cuda.current_device() else torch.cuda.device(idx)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            return self.ctx.__exit__(*exc)
        return False


_WS_BYTES = {}   # (n, dim, num_classes) -> workspace bytes


def _require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError("muscle_b200: `%s` must be a CUDA tensor (there is no CPU path)" % name)


def _as_feature_operand(x: torch.Tensor):
    """-> (contiguous tensor, layout, batch, pixels_per_image, dim)."""
    if x.dim() == 4:            # [B, D, h, w]
        b, d, h, w = x.shape
        return x.contiguous(), _lib.LAYOUT_NCHW, b, h * w, d
    if x.dim() == 3:            # [B, D, P]
        b, d, p = x.shape
        return x.contiguous(), _lib.LAYOUT_NCHW, b, p, d
    if x.dim() == 2:            # [N, D]
        n, d = x.shape
        return x.contiguous(), _lib.LAYOUT_PIXEL_MAJOR, 1, n, d
    raise ValueError("expected [B,D,h,w], [B,D,P] or [N,D] embeddings, got shape %s" % (tuple(x.shape),))


class _DenseInfoNCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cls, temperature, num_classes, eps, proto=None):
        _require_cuda(x, "x")
        _require_cuda(cls, "cls")
        if x.dtype != torch.float32:
            raise TypeError("embeddings must be float32 (the reference path is fp32)")
        lib = _lib.load()
        xc, layout, batch, ppi, dim = _as_feature_operand(x)
        n = batch * ppi
        cls_c = cls.reshape(-1).contiguous()
        if cls_c.numel() != n:
            raise ValueError("cls has %d entries for %d pixels" % (cls_c.numel(), n))
        if cls_c.dtype == torch.int64:
            cls64 = 1
        elif cls_c.dtype == torch.int32:
            cls64 = 0
        else:
            cls_c = cls_c.to(torch.int64)
            cls64 = 1
        need_grad = x.requires_grad or (proto is not None and proto.requires_grad)
        proto_c = None
        if proto is not None:
            _require_cuda(proto, "proto")
            if tuple(proto.shape) != (int(num_classes), dim):
                raise ValueError("proto must be [num_classes, D] = [%d, %d]" % (num_classes, dim))
            proto_c = proto.detach().contiguous().float()
        key = (n, dim, int(num_classes))
        ws_bytes = _WS_BYTES.get(key)
        if ws_bytes is None:
            ws_bytes = _WS_BYTES[key] = lib.muscle_dense_infonce_workspace_bytes(n, dim, num_classes)
        if ws_bytes == 0:
            raise ValueError("unsupported shape: dim=%d (64/128/256), num_classes=%d (<=%d)"
                             % (dim, num_classes, _lib.MAX_CLASSES))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        loss = torch.empty((), dtype=torch.float32, device=x.device)
        # forward: prepare -> plan -> row sums -> loss; the gradient pass runs in backward() from the
        # state kept in `ws`, with the upstream gradient multiplied in by the last kernel
        with _on_device(x.device):
            rc = lib.muscle_dense_infonce_proto_fwd(
                xc.data_ptr(), layout, batch, ppi, dim, cls_c.data_ptr(), cls64, num_classes,
                float(temperature), float(eps), None if proto_c is None else proto_c.data_ptr(), loss.data_ptr(),
                ws.data_ptr(), ws_bytes, _stream_ptr(x.device))
        _lib.check(rc, "dense_infonce")
        ctx.has_proto = proto_c is not None
        if need_grad:
            ctx.ws, ctx.ws_bytes = ws, ws_bytes
            ctx.args = (layout, batch, ppi, dim, int(num_classes), float(temperature))
            ctx.x_shape, ctx.xc_shape = x.shape, xc.shape
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        ws = getattr(ctx, "ws", None)
        if ws is None:
            return None, None, None, None, None, None
        lib = _lib.load()
        layout, batch, ppi, dim, num_classes, temperature = ctx.args
        g = grad_out.detach()
        if g.dtype != torch.float32 or g.device != ws.device or not g.is_contiguous():
            g = g.to(device=ws.device, dtype=torch.float32).contiguous()
        dx = torch.empty(ctx.xc_shape, dtype=torch.float32, device=ws.device)
        dproto = torch.empty((num_classes, dim), dtype=torch.float32, device=ws.device) if ctx.has_proto else None
        with _on_device(ws.device):
            rc = lib.muscle_dense_infonce_proto_bwd(layout, batch, ppi, dim, num_classes, temperature, g.data_ptr(),
                                                    1 if ctx.has_proto else 0, dx.data_ptr(),
                                                    None if dproto is None else dproto.data_ptr(), ws.data_ptr(),
                                                    ctx.ws_bytes, _stream_ptr(ws.device))
        _lib.check(rc, "dense_infonce backward")
        ctx.ws = None   # the 8-slot partial buffer is the large part of the state: free it now
        return dx.reshape(ctx.x_shape), None, None, None, None, dproto


def dense_infonce(x, cls, temperature=0.1, num_classes=21, eps=1e-12, proto=None):
    """Dense label-masked pixel InfoNCE over a batch (SURVEY.md §8c composition of the reference
    ``info_nce``, src/loss_multilabel.py:9-21): every valid pixel is a query, its positive is the
    prototype (mean normalised embedding) of its class, its negatives are all pixels of other
    classes; mean over valid pixels.

    x   : float32 CUDA [B, D, h, w] (the model's layout), [B, D, P] or [N, D]; D in {64,128,256}
    cls : per-pixel class ids, same pixel order, int64/int32; < 0 = ignore
    proto : optional [num_classes, D] prototypes that replace the hard class means as the positives — e.g.
            CAM-weighted ones from :func:`muscle_b200.pool.pool_prototypes` (SURVEY.md §8c soft variant).  The
            gradient then reaches ``proto`` (and, through the pooling op, the pixels and the weights); no
            gradient flows to the pixels through the prototypes inside this op.
    Returns a 0-dim float32 tensor with grad to ``x`` (forward and input gradient are produced by
    one fused pass; backward only scales)."""
    return _DenseInfoNCE.apply(x, cls, temperature, num_classes, eps, proto)


def dense_infonce_value_and_grad(x, cls, temperature=0.1, num_classes=21, eps=1e-12):
    """Loss and its input gradient from ONE call of the fused C entry point
    (``muscle_dense_infonce_fwd_bwd``), without autograd bookkeeping: for a loss at the end of the
    graph, ``loss, dx = dense_infonce_value_and_grad(feat.detach(), cls); feat.backward(dx)`` saves
    the autograd engine's thread hand-off (~0.1 ms of host time per step at a 0.3 ms step).
    Same operands and checks as :func:`dense_infonce`; returns (0-dim loss, dx shaped like ``x``)."""
    _require_cuda(x, "x")
    _require_cuda(cls, "cls")
    if x.dtype != torch.float32:
        raise TypeError("embeddings must be float32 (the reference path is fp32)")
    lib = _lib.load()
    xc, layout, batch, ppi, dim = _as_feature_operand(x.detach())
    n = batch * ppi
    cls_c = cls.reshape(-1).contiguous()
    if cls_c.numel() != n:
        raise ValueError("cls has %d entries for %d pixels" % (cls_c.numel(), n))
    if cls_c.dtype not in (torch.int64, torch.int32):
        cls_c = cls_c.to(torch.int64)
    cls64 = 1 if cls_c.dtype == torch.int64 else 0
    key = (n, dim, int(num_classes))
    ws_bytes = _WS_BYTES.get(key)
    if ws_bytes is None:
        ws_bytes = _WS_BYTES[key] = lib.muscle_dense_infonce_workspace_bytes(n, dim, num_classes)
    if ws_bytes == 0:
        raise ValueError("unsupported shape: dim=%d (64/128/256), num_classes=%d (<=%d)"
                         % (dim, num_classes, _lib.MAX_CLASSES))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    loss = torch.empty((), dtype=torch.float32, device=x.device)
    dx = torch.empty(xc.shape, dtype=torch.float32, device=x.device)
    with _on_device(x.device):
        rc = lib.muscle_dense_infonce_fwd_bwd(
            xc.data_ptr(), layout, batch, ppi, dim, cls_c.data_ptr(), cls64, num_classes,
            float(temperature), float(eps), loss.data_ptr(), dx.data_ptr(), ws.data_ptr(), ws_bytes,
            _stream_ptr(x.device))
    _lib.check(rc, "dense_infonce_value_and_grad")
    return loss, dx.reshape(x.shape)


def debug_tile(a_bf16: torch.Tensor, b_bf16: torch.Tensor):
    """S = A B^T [128,128] and G = bf16(S) B [128,D] through the tcgen05 building blocks: a test fixture, served by
    libmuscle_b200_diag.so (include/muscle_b200_diag.h), not by the product library."""
    lib = _lib.load_diag()
    assert a_bf16.dtype == torch.bfloat16 and b_bf16.dtype == torch.bfloat16
    assert a_bf16.shape == b_bf16.shape and a_bf16.shape[0] == 128
    a = a_bf16.contiguous()
    b = b_bf16.contiguous()
    d = a.shape[1]
    s = torch.empty(128, 128, dtype=torch.float32, device=a.device)
    g = torch.empty(128, d, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        rc = lib.muscle_debug_tile(a.data_ptr(), b.data_ptr(), d, s.data_ptr(), g.data_ptr(), _stream_ptr(a.device))
    if rc != 0:
        raise RuntimeError("muscle_debug_tile failed (code %d)" % rc)
    return s, g
