"""Drop-in for ``torchutils.get_dynamic_crops`` (src/torchutils.py:217-291; train_mcl.py:220).

The window plan is host integer work and consumes numpy's global RNG exactly like the reference
(:246-254, one image after the other); the tensor work — every bilinear resize and average pool
of the batch — is ONE kernel launch writing into a flat buffer.  The returned nested lists hold
views of that buffer, so any consumer works; ``muscle_b200.loss_multilabel.EMD`` recognises the
``CropPack`` lists and reads the flat buffer directly."""
from __future__ import annotations

import numpy as np
import torch

from ._call import call, require_cuda_f32


def _plan_image(h, w, rng):
    """Integer half of src/torchutils.py:240-280 for one h x w overlap box.  -> None (box rejected)
    or (view-1 windows [(top,left,lh,lw)], (rh,rw), pooled?, view-2 windows)."""
    if h < 15 or w < 15 or h / w > 5 or w / h > 5:                      # :240
        return None
    lh = rng.randint(h // 3, h // 2 + 1)                               # :246-247
    lw = rng.randint(w // 3, w // 2 + 1)
    while lh < 5 or lw < 5:                                            # :249-251
        lh = rng.randint(h // 3, h // 2 + 1)
        lw = rng.randint(w // 3, w // 2 + 1)
    sh = rng.randint(lh // 2, lh + 1)                                  # :253-254
    sw = rng.randint(lw // 2, lw + 1)
    rh, rw = round(h / (h / sh)), round(w / (w / sw))                  # :255-256,266-267
    wins1 = []
    if rh >= 7 and rw >= 7:                                            # :268
        for top in range(0, h, sh):
            for left in range(0, w, sw):
                if top + lh <= h and left + lw <= w:                   # :259
                    wins1.append((top, left, lh, lw))
    wins2 = [(top, left, min(h // 2, h - top), min(w // 2, w - left))
             for top in range(0, h - 1, h // 2) for left in range(0, w - 1, w // 2)]   # :276-278
    return wins1, (rh, rw), (rh > 28 or rw > 28), wins2


class CropPack(list):
    """A list of per-image lists of [1,C,h,w] crop tensors that also carries the flat buffer they
    are views of (``flat``), per-crop (offset, pixels) (``meta``, same nesting) and the channel count."""
    flat = None
    meta = None
    channels = 0


class _Crops(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, desc, offs, n_crops, total):
        xc = x.contiguous()
        b, c, h, w = xc.shape
        flat = torch.empty(total, dtype=torch.float32, device=x.device)
        call("muscle_crops_fwd", x.device, xc.data_ptr(), c, h, w, desc.data_ptr(), offs.data_ptr(), n_crops, total,
             flat.data_ptr())
        ctx.save_for_backward(desc, offs)
        ctx.dims = (b, c, h, w, n_crops, total)
        return flat

    @staticmethod
    def backward(ctx, dflat):
        desc, offs = ctx.saved_tensors
        b, c, h, w, n_crops, total = ctx.dims
        dx = torch.empty((b, c, h, w), dtype=torch.float32, device=dflat.device)
        call("muscle_crops_bwd", dflat.device, dflat.contiguous().data_ptr(), b, c, h, w, desc.data_ptr(),
             offs.data_ptr(), n_crops, total, dx.data_ptr())
        return dx, None, None, None, None


def _extract(x, rows, shapes_nested):
    """rows: list of 10-int descriptors (flat order); shapes_nested: per image list of (oh, ow).
    -> CropPack"""
    c = x.shape[1]
    offs, total = [], 0
    for r in rows:
        offs.append(total)
        total += c * r[8] * r[9]
    offs.append(total)
    dev = x.device
    desc_t = torch.tensor(rows, dtype=torch.int32).to(dev, non_blocking=True)
    offs_t = torch.tensor(offs, dtype=torch.int64).to(dev, non_blocking=True)
    flat = _Crops.apply(x, desc_t, offs_t, len(rows), total)
    pack = CropPack()
    pack.flat, pack.channels, pack.meta = flat, c, []
    i = 0
    for shapes in shapes_nested:
        img, meta = [], []
        for oh, ow in shapes:
            img.append(flat[offs[i]:offs[i + 1]].view(1, c, oh, ow))
            meta.append((offs[i], oh * ow))
            i += 1
        pack.append(img)
        pack.meta.append(meta)
    return pack


def get_dynamic_crops(x1, coord1, x2, coord2):
    """src/torchutils.py:217-291.  x1, x2 [B,C,H,W] fp32 CUDA; coord [B,4] = (top, left, h, w).
    -> (crops1, crops2, batch_indices): nested lists of [1,C,h',w'] tensors (grad flows to x1/x2)."""
    require_cuda_f32(x1, "x1")
    require_cuda_f32(x2, "x2")
    bsz, _, H, W = x1.shape
    c1 = coord1.detach().to("cpu").long().tolist()
    c2 = coord2.detach().to("cpu").long().tolist()
    rows1, rows2, shapes1, shapes2, batch_indices = [], [], [], [], []
    for b in range(bsz):
        t1, l1, bh, bw = c1[b]
        t2, l2, bh2, bw2 = c2[b]
        h, w = min(bh, H - t1), min(bw, W - l1)          # python slicing clamps (:226)
        h2, w2 = min(bh2, H - t2), min(bw2, W - l2)
        plan = _plan_image(h, w, np.random)
        if plan is None:
            continue
        wins1, (rh, rw), pool, wins2 = plan
        if not wins1:
            continue
        r1, s1 = [], []
        for top, left, lh, lw in wins1:
            oh, ow = (rh // 4, rw // 4) if pool else (rh, rw)
            r1.append([b, t1 + top, l1 + left, lh, lw, rh, rw, int(pool), oh, ow])
            s1.append((oh, ow))
        r2, s2 = [], []
        for top, left, hh, ww in wins2:
            hh, ww = min(hh, h2 - top), min(ww, w2 - left)   # slicing ins2 clamps at its own size
            r2.append([b, t2 + top, l2 + left, hh, ww, hh, ww, 1, hh // 4, ww // 4])
            s2.append((hh // 4, ww // 4))
        rows1 += r1
        rows2 += r2
        shapes1.append(s1)
        shapes2.append(s2)
        batch_indices.append(b)
    if not rows1:
        return [], [], []
    return _extract(x1, rows1, shapes1), _extract(x2, rows2, shapes2), batch_indices


class CropList(list):
    """A flat list of [1,C,h,w] crop tensors that are views of ONE buffer (``flat``); ``offsets[i]`` is
    crop i's first element in it.  ``EMD.static_matching`` reads the buffer directly."""
    flat = None
    offsets = None


def _extract_flat(x, rows):
    """rows: 10-int window descriptors with resize target == window size (an exact copy: every
    bilinear tap has weight 1 or 0).  One gather launch forward, one scatter launch backward —
    python slicing would cost one autograd node and one full-size zero-filled gradient per window."""
    c = x.shape[1]
    offs, total = [], 0
    for r in rows:
        offs.append(total)
        total += c * r[8] * r[9]
    offs.append(total)
    desc_t = torch.tensor(rows, dtype=torch.int32).to(x.device, non_blocking=True)
    offs_t = torch.tensor(offs, dtype=torch.int64).to(x.device, non_blocking=True)
    flat = _Crops.apply(x, desc_t, offs_t, len(rows), total)
    out = CropList(flat[offs[i]:offs[i + 1]].view(1, c, r[8], r[9]) for i, r in enumerate(rows))
    out.flat, out.offsets = flat, offs[:-1]
    return out


def get_static_crops(x1, coord1, x2, coord2):
    """src/torchutils.py:177-214.  x1, x2 [B,C,H,W] fp32 CUDA; coord [B,4] = (top, left, h, w).
    -> (crops1, crops2, batch_indices): flat lists of nine half-size windows of view 1's overlap box
    and nine third-size windows of view 2's per kept image, as [1,C,·,·] tensors (grad flows to
    x1/x2); boxes under 15 px or more elongated than 7:1 are skipped (:199).  Window sizes follow
    python slicing (clamped at the box, :186,:191,:205,:212).  Feed to ``EMD()(crops1, crops2)``."""
    require_cuda_f32(x1, "x1")
    require_cuda_f32(x2, "x2")
    H1, W1 = x1.shape[2:]
    H2, W2 = x2.shape[2:]
    c1 = coord1.detach().to("cpu").long().tolist()
    c2 = coord2.detach().to("cpu").long().tolist()
    rows1, rows2, batch_indices = [], [], []

    def window(b, top, left, bh, bw, i, j, hh, ww):      # ins[:, :, i:i+hh, j:j+ww] of a bh x bw box
        hh, ww = min(hh, bh - i), min(ww, bw - j)
        if hh < 1 or ww < 1:
            raise ValueError("static crop window is empty (view-2 box smaller than view 1's window grid)")
        return [b, top + i, left + j, hh, ww, hh, ww, 0, hh, ww]

    for b in range(x1.shape[0]):
        t1, l1, bh1, bw1 = c1[b]
        t2, l2, bh2, bw2 = c2[b]
        h, w = max(min(bh1, H1 - t1), 0), max(min(bw1, W1 - l1), 0)      # both window grids use view 1's box size (:197)
        h2, w2 = max(min(bh2, H2 - t2), 0), max(min(bw2, W2 - l2), 0)
        if h < 15 or w < 15 or h / w > 7 or w / h > 7:
            continue
        for i in (0, h // 4, h // 2):
            for j in (0, w // 4, w // 2):
                rows1.append(window(b, t1, l1, h, w, i, j, h // 2, w // 2))
        for i in (0, h // 3, 2 * h // 3):
            for j in (0, w // 3, 2 * w // 3):
                rows2.append(window(b, t2, l2, h2, w2, i, j, h // 3, w // 3))
                batch_indices.append(b)
    if not rows1:
        return [], [], []
    return _extract_flat(x1, rows1), _extract_flat(x2, rows2), batch_indices
