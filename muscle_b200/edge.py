"""Drop-in for the reference's ``edge.FieldLoss`` (src/edge.py:175-384; constructed at
train_muscle.py:163, called at :196): same constructor and call signature, same return convention
``(loss | False | 0, mag_fg.sum(1))`` and the same consumption of Python's global ``random`` stream
(``random.sample`` of the outside list, then of the inside list, per qualifying (image, class) in
image-major order, src/edge.py:297-299).  The arithmetic runs in libmuscle_b200.so."""
from __future__ import annotations

import random

import torch
from torch import nn

from . import _lib
from ._call import call, ptr, require_cuda_f32


class _FieldPairLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dense_ft, mask, planes, offsets, sel, lists, n_active, k):
        n, ch, h, w = dense_ft.shape
        lib = _lib.load()
        dev = dense_ft.device
        d = dense_ft.contiguous()
        m = mask.contiguous()
        sb = lib.muscle_field_loss_scratch_bytes(n_active, k, ch, m.shape[1])
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dd = torch.empty_like(d) if dense_ft.requires_grad else None
        call("muscle_field_loss", dev, d.data_ptr(), m.data_ptr(), n, ch, m.shape[1], h, w, planes.data_ptr(),
             offsets.data_ptr(), sel.data_ptr(), lists.data_ptr(), n_active, k, loss.data_ptr(), ptr(dd),
             scratch.data_ptr(), sb)
        ctx.save_for_backward(dd)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        dd, = ctx.saved_tensors
        return (None if dd is None else dd * gloss), None, None, None, None, None, None, None


class _FieldPairLossDecoder(torch.autograd.Function):
    """stage 3 with dense_ft interpolated on the fly from the decoder output (csrc/fieldloss.cu: muscle_field_loss_decoder)"""

    @staticmethod
    def forward(ctx, p3_dec, mask, planes, offsets, sel, lists, n_active, k, height, width):
        n, ch, hs, ws = p3_dec.shape
        lib = _lib.load()
        dev = p3_dec.device
        d = p3_dec.contiguous()
        m = mask.contiguous()
        sb = lib.muscle_field_loss_scratch_bytes(n_active, k, ch, m.shape[1])
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dd = torch.empty_like(d) if p3_dec.requires_grad else None
        call("muscle_field_loss_decoder", dev, d.data_ptr(), hs, ws, m.data_ptr(), n, ch, m.shape[1], height, width,
             planes.data_ptr(), offsets.data_ptr(), sel.data_ptr(), lists.data_ptr(), n_active, k, loss.data_ptr(), ptr(dd),
             scratch.data_ptr(), sb)
        ctx.save_for_backward(dd)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        dd, = ctx.saved_tensors
        return (None if dd is None else dd * gloss), None, None, None, None, None, None, None, None, None


class FieldLoss(nn.Module):
    """src/edge.py:175-193.  sobel_size 3 or 5 (src/edge.py:33-43; train_muscle.py:163 passes 5)."""

    def __init__(self, num_classes=21, gaussian_size=7, k=100, guassian_sigma=None, sobel_size=5, beta=1e3):
        super().__init__()
        if sobel_size not in (3, 5):
            raise ValueError("sobel_size must be 3 or 5 (src/edge.py:33-43 defines no other filter)")
        self.sobel_size = int(sobel_size)
        self.num_classes = num_classes
        self.num_fg_cls = num_classes - 1
        self.k = k
        self.beta = float(beta)

    def edges(self, seg_map, label_with_bg, step):
        """Stages a13-a15 on the device.  -> dict(mag_sum [n,H,W], orient u8 [n,C-1,H,W] (bins 0..7),
        pos u8 [n,C-1,H,W], counts int32 [n,C-1,3] = (#boundary, #outside kept, #inside kept))."""
        require_cuda_f32(seg_map, "seg_map")
        n, c, h, w = seg_map.shape
        dev = seg_map.device
        lab = label_with_bg.to(device=dev, dtype=torch.float32).contiguous()
        prob = torch.empty((n, c - 1, h, w), dtype=torch.float32, device=dev)
        mag = torch.empty_like(prob)
        orient = torch.empty((n, c - 1, h, w), dtype=torch.uint8, device=dev)
        pos = torch.empty_like(orient)
        mag_sum = torch.empty((n, h, w), dtype=torch.float32, device=dev)
        plane_max = torch.empty((n, c - 1), dtype=torch.int32, device=dev)
        counts = torch.empty((n, c - 1, 3), dtype=torch.int32, device=dev)
        call("muscle_field_edges_sobel", dev, seg_map.detach().contiguous().data_ptr(), lab.data_ptr(), n, c, h, w,
             self.beta, int(step), self.sobel_size, prob.data_ptr(), mag.data_ptr(), orient.data_ptr(), pos.data_ptr(),
             mag_sum.data_ptr(), plane_max.data_ptr(), counts.data_ptr())
        return dict(mag_sum=mag_sum, orient=orient, pos=pos, counts=counts, label=lab, mag=mag)

    def _select(self, st, n, c, h, w, step, dev):
        """Host half of the forward (src/edge.py:378, :297-299): gates and Python's random stream.  -> None (gate failed:
        the caller returns False), () (nothing contributed: python 0) or (planes_t, offsets_t, sel_t, lists, n_active)."""
        counts = st["counts"].cpu()            # the one device->host read (the reference syncs per class)
        label = st["label"].cpu()
        if int(counts[:, :, 0].sum()) < 10:    # src/edge.py:377-378
            return None
        k = self.k
        planes, offsets, sel, total = [], [], [], 0
        for b in range(n):
            for cc in range(c - 1):
                if label[b, 1 + cc] == 0:
                    continue
                n_out, n_in = int(counts[b, cc, 1]), int(counts[b, cc, 2])
                if n_in > k and n_out > k:                       # :297
                    rand_out = random.sample(range(n_out), k)    # :298  (outside first)
                    rand_in = random.sample(range(n_in), k)      # :299
                    planes.append(b * (c - 1) + cc)
                    offsets += [total, total + n_out]
                    total += n_out + n_in
                    sel += rand_out + rand_in
        if not planes:
            return ()
        planes_t = torch.tensor(planes, dtype=torch.int32).to(dev, non_blocking=True)
        offsets_t = torch.tensor(offsets, dtype=torch.int64).to(dev, non_blocking=True)
        sel_t = torch.tensor(sel, dtype=torch.int32).to(dev, non_blocking=True)
        lists = torch.empty(total, dtype=torch.int32, device=dev)
        call("muscle_field_compact", dev, st["orient"].data_ptr(), st["pos"].data_ptr(), h, w, int(step),
             planes_t.data_ptr(), offsets_t.data_ptr(), len(planes), lists.data_ptr())
        self._last = dict(planes=planes, offsets=offsets, lists=lists)   # for tests / inspection
        return planes_t, offsets_t, sel_t, lists, len(planes)

    def forward(self, seg_map, dense_ft, mask, label_with_bg, step=7):
        require_cuda_f32(dense_ft, "dense_ft")
        n, c, h, w = seg_map.shape
        if dense_ft.shape[0] != n or dense_ft.shape[2:] != seg_map.shape[2:]:
            raise ValueError("seg_map and dense_ft must share batch and spatial size")
        st = self.edges(seg_map, label_with_bg, step)
        mag_sum = st["mag_sum"]
        picked = self._select(st, n, c, h, w, step, seg_map.device)
        if picked is None:
            return False, mag_sum
        if not picked:
            return 0, mag_sum                  # :359-361 (python 0: nothing contributed)
        planes_t, offsets_t, sel_t, lists, n_active = picked
        mask_f = mask.detach().to(device=seg_map.device, dtype=torch.float32)
        loss = _FieldPairLoss.apply(dense_ft, mask_f, planes_t, offsets_t, sel_t, lists, n_active, self.k)
        return loss, mag_sum

    def forward_from_decoder(self, p3_dec, fuse_dec_weight, fuse_dec_bias, mask, label_with_bg, step=7, size=None):
        """The decoder tail of src/MuSCLe.py:285-286 and this loss in one: what the trainer gets
        from ``model(img, cam='seg')`` followed by ``criterion2(seg_map, dense_ft, mask, label_with_bg, step)``
        (train_muscle.py:188-196), from the DECODER OUTPUT ``p3_dec`` [B, ch, H/8, W/8] and ``fuse_dec``'s parameters.
        ``dense_ft`` ([B, ch, H, W]: 4.3 GB at batch 16, 512x512, 256 channels) is never formed: the 1x1 ``fuse_dec``
        runs at the decoder's resolution (it commutes with the bilinear up-sampling), only its 21 channels are
        up-sampled for the edge stage, and the <= 2k sampled columns per plane are interpolated from ``p3_dec`` on the
        fly.  Returns ``(loss | False | 0, mag_fg.sum(1), seg_map)``; the gradient reaches ``p3_dec`` (``seg_map`` enters the
        loss without gradient in the reference as well, src/edge.py:366-374)."""
        from .er import upsample_bilinear
        require_cuda_f32(p3_dec, "p3_dec")
        n, ch, hs, ws = p3_dec.shape
        h, w = (int(size[0]), int(size[1])) if size is not None else (mask.shape[2], mask.shape[3])
        dev = p3_dec.device
        wt = fuse_dec_weight.detach().reshape(fuse_dec_weight.shape[0], -1).contiguous().float()
        c = wt.shape[0]
        with torch.no_grad():
            seg_small = torch.empty((n, c, hs, ws), dtype=torch.float32, device=dev)
            bias = None if fuse_dec_bias is None else fuse_dec_bias.detach().contiguous().float()
            call("muscle_conv1x1_fwd", dev, p3_dec.detach().contiguous().data_ptr(), wt.data_ptr(), ptr(bias), n, ch, c,
                 hs * ws, seg_small.data_ptr())
            seg_map = upsample_bilinear(seg_small, (h, w))
        st = self.edges(seg_map, label_with_bg, step)
        mag_sum = st["mag_sum"]
        picked = self._select(st, n, c, h, w, step, dev)
        if picked is None:
            return False, mag_sum, seg_map
        if not picked:
            return 0, mag_sum, seg_map
        planes_t, offsets_t, sel_t, lists, n_active = picked
        mask_f = mask.detach().to(device=dev, dtype=torch.float32)
        loss = _FieldPairLossDecoder.apply(p3_dec, mask_f, planes_t, offsets_t, sel_t, lists, n_active, self.k, h, w)
        return loss, mag_sum, seg_map
