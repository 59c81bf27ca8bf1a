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


def image_level_contrast_nosync(emb, label):
    """The same loss without the device->host read: -> (loss, n_used), both 0-dim CUDA tensors.  `loss` is 0 (with a
    zero gradient) when `n_used` == 0, i.e. where the reference returns python ``0``; a training loop can add it
    unconditionally and never stalls the stream (SURVEY.md 7.3 asks for this form beside the reference's convention)."""
    require_cuda_f32(emb, "emb")
    if not label.is_cuda:
        raise RuntimeError("muscle_b200: `label` must be a CUDA tensor (there is no CPU path)")
    return _ImageLevelContrast.apply(emb, label)


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


class _EmdDynamic(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flat1, flat2, pair_desc, n_pairs, n_images, channels, max_n, max_m, max_nm, reg, maxiter):
        lib = _lib.load()
        dev = flat1.device
        f1, f2 = flat1.contiguous(), flat2.contiguous()
        sb = lib.muscle_emd_scratch_bytes(n_pairs, n_images, max_nm, maxiter)
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        d1 = torch.empty_like(f1) if flat1.requires_grad else None
        call("muscle_emd_dynamic_fwd_bwd", dev, f1.data_ptr(), f2.data_ptr(), pair_desc.data_ptr(), n_pairs, n_images,
             channels, max_n, max_m, max_nm, float(reg), int(maxiter), loss.data_ptr(), ptr(d1), f1.numel(),
             scratch.data_ptr(), sb)
        ctx.save_for_backward(d1)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        d1, = ctx.saved_tensors
        return (None if d1 is None else d1 * gloss,) + (None,) * 10


class _EmdStatic(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flat1, flat2, pair_desc, n_pairs, max_planes):
        lib = _lib.load()
        dev = flat1.device
        f1, f2 = flat1.contiguous(), flat2.contiguous()
        sb = lib.muscle_emd_static_scratch_bytes(n_pairs, max_planes)
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dists = torch.empty(n_pairs, dtype=torch.float32, device=dev)
        d1 = torch.empty_like(f1) if flat1.requires_grad else None
        call("muscle_emd_static_fwd_bwd", dev, f1.data_ptr(), f2.data_ptr(), pair_desc.data_ptr(), n_pairs, max_planes,
             loss.data_ptr(), dists.data_ptr(), ptr(d1), f1.numel(), scratch.data_ptr(), sb)
        ctx.save_for_backward(d1)
        ctx.mark_non_differentiable(dists)
        return loss, dists

    @staticmethod
    def backward(ctx, gloss, _gdists):
        d1, = ctx.saved_tensors
        return (None if d1 is None else d1 * gloss,) + (None,) * 4


def _static_target_shape(h, w, long_side=7):
    """src/loss_multilabel.py:265-269, :275-278 (python round: half to even, as the reference)."""
    if w < h:
        return int(round(w * long_side / h)), long_side
    return long_side, int(round(h * long_side / w))


def _pack_regions(crops):
    """nested lists of [1,C,h,w] tensors -> (flat, meta, C).  Lists made by
    muscle_b200.torchutils.get_dynamic_crops already carry their flat buffer; anything else is
    concatenated (torch.cat is plumbing: autograd routes the gradient back to every tensor)."""
    from .torchutils import CropPack
    if isinstance(crops, CropPack) and crops.flat is not None:
        return crops.flat, crops.meta, crops.channels
    pieces, meta, off, ch = [], [], 0, None
    for img in crops:
        m = []
        for t in img:
            require_cuda_f32(t, "crop")
            if t.dim() != 4 or t.shape[0] != 1:
                raise ValueError("crops must be [1,C,h,w] tensors")
            ch = t.shape[1] if ch is None else ch
            pieces.append(t.reshape(-1))
            m.append((off, t.shape[2] * t.shape[3]))
            off += t.numel()
        meta.append(m)
    return torch.cat(pieces), meta, ch


class EMD:
    """src/loss_multilabel.py:108-338: ``EMD()(crops1, crops2, mode='dynamic')`` (train_mcl.py:147,221,
    the trainer's path) and ``mode='static'`` (:259-285, the class default, fed by get_static_crops).  Gradient flows to crops1; crops2 and the Sinkhorn
    marginals are constants, as in the reference (:308-313, train_mcl.py:220 passes vw2.detach())."""

    reg = 0.1        # :207
    maxiter = 10     # :313,323

    def dynamic_matching(self, crops1, crops2):
        if len(crops1) == 0 or len(crops1) != len(crops2):
            raise ValueError("crops1 and crops2 must be non-empty lists of equal length")
        flat1, meta1, ch = _pack_regions(crops1)
        flat2, meta2, _ = _pack_regions(crops2)
        if flat2.requires_grad:
            flat2 = flat2.detach()      # region 2 carries no gradient on the reference's path
        rows, max_n, max_m, max_nm = [], 0, 0, 0
        for i, (m1, m2) in enumerate(zip(meta1, meta2)):
            for off1, n in m1:                      # crop1 outer, crop2 inner (:295,302)
                for off2, m in m2:
                    rows.append([off1, n, off2, m, i, 0])
                    max_n, max_m, max_nm = max(max_n, n), max(max_m, m), max(max_nm, n + m)
        desc = torch.tensor(rows, dtype=torch.int64).to(flat1.device, non_blocking=True)
        return _EmdDynamic.apply(flat1, flat2, desc, len(rows), len(meta1), ch, max_n, max_m, max_nm, self.reg,
                                 self.maxiter)

    def static_matching(self, crops1, crops2, return_dists=False):
        """src/loss_multilabel.py:259-285.  crops1, crops2: flat lists of [n,C,h,w] tensors of equal
        length >= 2 (what get_static_crops makes); pair i is (crops1[i], crops2[i]).  Returns the sum
        of the two smallest pair distances (0-dim tensor; grad flows to crops1 only, :273)."""
        if len(crops1) != len(crops2) or len(crops1) < 2:
            raise ValueError("static matching needs two lists of equal length >= 2 (the reference sums dists[0] + dists[1])")
        from .torchutils import CropList
        rows, p1, p2, off1, off2, max_planes = [], [], [], 0, 0, 0

        def is_packed(crops):     # still the untouched list get_static_crops made (same tensors, same order)?
            if not (isinstance(crops, CropList) and crops.flat is not None and len(crops.offsets) == len(crops)):
                return False
            base = crops.flat.data_ptr()
            return all(torch.is_tensor(t) and t.data_ptr() == base + 4 * o for t, o in zip(crops, crops.offsets))
        packed1, packed2 = is_packed(crops1), is_packed(crops2)
        for k, (a, b) in enumerate(zip(crops1, crops2)):
            require_cuda_f32(a, "crop")
            require_cuda_f32(b, "crop")
            if a.dim() != 4 or b.dim() != 4 or a.shape[:2] != b.shape[:2]:
                raise ValueError("paired crops must be [n,C,h,w] tensors with the same n and C")
            n, c, h1, w1 = a.shape
            h2, w2 = b.shape[2:]
            (oh1, ow1), (oh2, ow2) = _static_target_shape(h1, w1), _static_target_shape(h2, w2)
            if min(h1, w1, h2, w2, oh1, ow1, oh2, ow2) < 1 or max(h1, w1, h2, w2) > 1024:
                raise ValueError("crop sides must be 1..1024 px and resize to at least one pixel (got %s, %s)"
                                 % (tuple(a.shape), tuple(b.shape)))
            if packed1:
                off1 = crops1.offsets[k]
            if packed2:
                off2 = crops2.offsets[k]
            rows.append([off1, n * c, h1, w1, oh1, ow1, off2, h2, w2, oh2, ow2, n])
            if not packed1:
                p1.append(a.reshape(-1))
                off1 += a.numel()
            if not packed2:
                p2.append(b.detach().reshape(-1))
                off2 += b.numel()
            max_planes = max(max_planes, n * c)
        # lists made by torchutils.get_static_crops already are views of one buffer; anything else is
        # concatenated (plumbing: autograd routes d flat1 back to every crop)
        flat1 = crops1.flat if packed1 else torch.cat(p1)
        flat2 = crops2.flat.detach() if packed2 else torch.cat(p2)
        desc = torch.tensor(rows, dtype=torch.int64).to(flat1.device, non_blocking=True)
        loss, dists = _EmdStatic.apply(flat1, flat2, desc, len(rows), max_planes)
        return (loss, dists) if return_dists else loss

    def __call__(self, crops1, crops2, mode="static"):
        if mode == "dynamic":
            return self.dynamic_matching(crops1, crops2)
        if mode == "static":
            return self.static_matching(crops1, crops2)
        print("Unrecognised mode!")      # :335-336
        return None


class _InfoNce(torch.autograd.Function):
    """mean / sum: value and the three gradients in ONE C call at forward time (backward multiplies by the upstream
    scalar).  none: forward = the [Q] row losses; backward calls the C path again with the upstream gradient as per-row
    weights (muscle_info_nce_fwd_bwd_ex)."""

    @staticmethod
    def _run(q, p, k, temperature, mode_mean, rq, rk, row_weight, want_rows, need):
        lib = _lib.load()
        dev = q.device
        nq, npos, nk, dim_pad = q.shape[0], p.shape[0], k.shape[0], q.shape[1]
        wsb = lib.muscle_info_nce_workspace_bytes(nq, nk, dim_pad)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        rows = torch.empty(nq, dtype=torch.float32, device=dev) if want_rows else None
        dq = torch.empty_like(q) if need[0] else None
        dp = torch.empty_like(p) if need[1] else None
        dk = torch.empty_like(k) if need[2] else None
        call("muscle_info_nce_fwd_bwd_ex", dev, q.data_ptr(), p.data_ptr(), k.data_ptr(), nq, npos, nk, dim_pad,
             float(temperature), 1 if mode_mean else 0, float(rq), float(rk), ptr(row_weight), loss.data_ptr(), ptr(rows),
             ptr(dq), ptr(dp), ptr(dk), ws.data_ptr(), wsb)
        return loss, rows, dq, dp, dk

    @staticmethod
    def forward(ctx, query, positive_keys, negative_keys, temperature, mode, dim_pad, rq, rk):
        def padded(t):      # zero columns change no dot product; the tensor-pipe passes need dim in {64,128,256}
            t = t.contiguous()
            return t if t.shape[1] == dim_pad else torch.nn.functional.pad(t, (0, dim_pad - t.shape[1]))

        q, p, k = padded(query), padded(positive_keys), padded(negative_keys)
        need = ctx.needs_input_grad[:3]
        ctx.mode, ctx.d = mode, query.shape[1]
        if mode == "none":
            _, rows, _, _, _ = _InfoNce._run(q, p, k, temperature, False, rq, rk, None, True, (False, False, False))
            ctx.save_for_backward(q, p, k)
            ctx.args = (float(temperature), float(rq), float(rk), tuple(need))
            return rows
        loss, _, dq, dp, dk = _InfoNce._run(q, p, k, temperature, mode == "mean", rq, rk, None, False, need)
        d = ctx.d
        ctx.save_for_backward(*(None if t is None else t[:, :d] for t in (dq, dp, dk)))
        return loss

    @staticmethod
    def backward(ctx, g):
        d = ctx.d
        if ctx.mode == "none":
            q, p, k = ctx.saved_tensors
            temperature, rq, rk, need = ctx.args
            gw = g.detach().to(torch.float32).contiguous()
            _, _, dq, dp, dk = _InfoNce._run(q, p, k, temperature, False, rq, rk, gw, False, need)
            return (None if dq is None else dq[:, :d], None if dp is None else dp[:, :d], None if dk is None else dk[:, :d],
                    None, None, None, None, None)
        dq, dp, dk = ctx.saved_tensors
        return (None if dq is None else dq * g, None if dp is None else dp * g, None if dk is None else dk * g,
                None, None, None, None, None)


def info_nce(query, positive_keys, negative_keys, temperature=0.1, reduction="mean"):
    """src/loss_multilabel.py:9-21.  query [Q,D], positive_keys [P,D], negative_keys [K,D], fp32 CUDA, rows of any norm
    (the reference does not normalise and neither does this: the largest row norms are read once -- one device->host
    copy of two floats -- and handed to the C path as uniform scales, see muscle_info_nce_fwd_bwd_ex).
    reduction in {'mean','sum','none'}, all three differentiable."""
    for name, t in (("query", query), ("positive_keys", positive_keys), ("negative_keys", negative_keys)):
        require_cuda_f32(t, name)
        if t.dim() != 2 or t.shape[1] != query.shape[1]:
            raise ValueError("info_nce operands must be [rows, D] with a common D")
    if reduction not in ("mean", "sum", "none"):
        raise ValueError("reduction must be 'mean', 'sum' or 'none'")
    d = query.shape[1]
    if d > 256:
        raise ValueError("info_nce supports D <= 256 (got %d)" % d)
    with torch.no_grad():
        rq = query.norm(dim=1).max()
        rk = torch.maximum(negative_keys.norm(dim=1).max(), positive_keys.mean(0).norm())
        rq, rk = (float(v) for v in torch.stack([rq, rk]).cpu())
    if not (rq > 0.0 and rk > 0.0 and rq < float("inf") and rk < float("inf")):
        raise ValueError("info_nce needs finite, non-zero queries and keys (max norms %g, %g)" % (rq, rk))
    dim_pad = 64 if d <= 64 else (128 if d <= 128 else 256)
    return _InfoNce.apply(query, positive_keys, negative_keys, temperature, reduction, dim_pad, rq, rk)


# --------------------------------------------------------------------------------------------------
# f4: the classification losses of the MCL stage
# --------------------------------------------------------------------------------------------------

class _ClsLosses(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, label, offset, gamma, alpha):
        lg = logits.contiguous()
        lb = label.contiguous().float()
        b, stride = lg.shape
        c = lb.shape[1]
        out = torch.empty(4, dtype=torch.float32, device=lg.device)
        ctx.save_for_backward(lg, lb)
        ctx.args = (b, c, stride, int(offset), float(gamma), float(alpha))
        call("muscle_cls_losses_fwd_bwd", lg.device, lg.data_ptr(), lb.data_ptr(), b, c, stride, int(offset), float(gamma),
             float(alpha), None, out.data_ptr(), None)
        return out

    @staticmethod
    def backward(ctx, gout):
        lg, lb = ctx.saved_tensors
        b, c, stride, offset, gamma, alpha = ctx.args
        # the four outputs are {total, focal, soft margin, pair}: the usual upstream is on the total alone
        g = gout.detach().float()
        if bool((g[1:] != 0).any()):
            raise NotImplementedError("differentiate loss_cls (output 0); the three parts are reported for logging")
        scale = g[0:1].contiguous()
        out = torch.empty(4, dtype=torch.float32, device=lg.device)
        dlg = torch.empty_like(lg)
        call("muscle_cls_losses_fwd_bwd", lg.device, lg.data_ptr(), lb.data_ptr(), b, c, stride, offset, gamma, alpha,
             scale.data_ptr(), out.data_ptr(), dlg.data_ptr())
        return dlg, None, None, None, None


def mcl_cls_loss(logits, label, gamma=2.0, alpha=0.5):
    """train_mcl.py:180-183 in one launch: ``FocalLoss()(sigmoid(l), y) + MultiLabelSoftMarginLoss()(l, y) +
    Log_Sum_Exp_Pairwise_Loss(sigmoid(l), y).mean()`` with l = logits[:, 1:] ([B, 21] logits incl. the background
    column) or logits itself ([B, 20]).  -> (loss_cls, loss_focal, loss_softmargin, loss_pair); differentiate loss_cls."""
    require_cuda_f32(logits, "logits")
    if logits.dim() != 2 or label.dim() != 2 or logits.shape[0] != label.shape[0]:
        raise ValueError("expected logits [B, C(+1)] and label [B, C]")
    offset = logits.shape[1] - label.shape[1]
    if offset not in (0, 1):
        raise ValueError("logits must have as many columns as the labels, or one more (background first)")
    out = _ClsLosses.apply(logits, label.to(logits.device), offset, gamma, alpha)
    return out[0], out[1], out[2], out[3]


class _Focal(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prob, target, gamma, alpha):
        p, t = prob.contiguous(), target.contiguous().float()
        b, c = p.shape
        out = torch.empty(4, dtype=torch.float32, device=p.device)
        dp = torch.empty_like(p) if prob.requires_grad else None
        call("muscle_focal_loss_fwd_bwd", p.device, p.data_ptr(), t.data_ptr(), b, c, float(gamma), float(alpha), None,
             out.data_ptr(), ptr(dp))
        ctx.save_for_backward(dp)
        return out[1]

    @staticmethod
    def backward(ctx, g):
        dp, = ctx.saved_tensors
        return (None if dp is None else dp * g), None, None, None


class FocalLoss(torch.nn.Module):
    """src/loss_multilabel.py:68-91: input = probabilities [n, c], target [n, c]; sum over classes, mean over images."""

    def __init__(self, gamma=2, alpha=0.5, size_average=True, weight=None):
        super().__init__()
        self.gamma, self.alpha, self.size_average, self.weight = gamma, alpha, size_average, weight

    def forward(self, input, target):
        require_cuda_f32(input, "input")
        return _Focal.apply(input, target.to(input.device), self.gamma, self.alpha)


class _Lsep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, labels):
        p, l = pred.contiguous(), labels.contiguous().float()
        b, c = p.shape
        rows = torch.empty(b, dtype=torch.float32, device=p.device)
        call("muscle_lsep_fwd_bwd", p.device, p.data_ptr(), l.data_ptr(), b, c, None, rows.data_ptr(), None)
        ctx.save_for_backward(p, l)
        return rows

    @staticmethod
    def backward(ctx, grows):
        p, l = ctx.saved_tensors
        b, c = p.shape
        g = grows.detach().contiguous().float()
        dp = torch.empty_like(p)
        call("muscle_lsep_fwd_bwd", p.device, p.data_ptr(), l.data_ptr(), b, c, g.data_ptr(), None, dp.data_ptr())
        return dp, None


def Log_Sum_Exp_Pairwise_Loss(pred, labels):
    """src/loss_multilabel.py:24-33: pred, labels [N, C] -> [N] (the caller takes the mean, train_mcl.py:182)."""
    require_cuda_f32(pred, "pred")
    return _Lsep.apply(pred, labels.to(pred.device))
