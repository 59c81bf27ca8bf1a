#!/usr/bin/env python
"""Device-time table of every op on the path besides the dense InfoNCE (which bench.py measures):
CUDA-event timing on the launching stream, L2 flushed between iterations, fwd+bwd through the
public Python mirrors; beside each, the oracle (the CPU restatement of the reference, fp32 torch
on all host cores) timed on the same inputs as the reported CPU baseline (bounded: a few
repeats).  Prints one JSON object; profiles/ keeps a copy per round.

    python scripts/bench_ops.py [--iters 20]
"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.nn.functional as F


def collect(iters=20, with_cpu=True, big=True):
    """-> dict of op records (device ms fwd+bwd, roofline-style figures; the oracle's CPU time beside each when
    `with_cpu`).  `big`: include the config-3-size FieldLoss entries (8.6 GB of device memory)."""
    class _A:
        pass
    args = _A()
    args.iters = iters
    from muscle_b200 import heads
    from muscle_b200.edge import FieldLoss
    from muscle_b200.loss_multilabel import EMD, PixPro, image_level_contrast
    from muscle_b200.torchutils import get_dynamic_crops
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}

    def timed(fn):
        for _ in range(3):
            fn()
        evs = []
        for _ in range(args.iters):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        return sum(x.elapsed_time(y) for x, y in evs) / len(evs)

    import time
    import workloads
    O = None
    if with_cpu:
        from oracle import muscle_oracle as O   # CPU baseline legs only (the checker is never on the product arm's path)

    def cpu_timed(fn, budget_s=6.0, max_reps=5):
        fn()
        t, reps = 0.0, 0
        while t < budget_s and reps < max_reps:
            t0 = time.perf_counter(); fn(); t += time.perf_counter() - t0; reps += 1
        return 1e3 * t / reps

    def beside(entry, cpu_fn, **kw):
        if not with_cpu:
            return entry
        cpu_ms = cpu_timed(cpu_fn, **kw)
        entry["cpu_ms"] = cpu_ms
        entry["cpu_cores"] = torch.get_num_threads()
        entry["cpu_kind"] = "port (oracle, fp32 torch CPU, same inputs)"
        entry["gpu_over_cpu"] = cpu_ms / entry["ms"]
        return entry

    def cpu_leaf(t):
        return t.detach().cpu().requires_grad_()

    out = {}
    # ---- PCM + CAM head, config 2 at stride 8 (B=8, 56x56, b7: Cin=336, p7 ch=640)
    B, P_h = 8, 56
    cam = torch.rand(B, 21, P_h, P_h, generator=gen).to(dev).requires_grad_()
    f = torch.relu(torch.randn(B, 336, P_h, P_h, generator=gen)).to(dev)
    wt = (torch.randn(128, 336, 1, 1, generator=gen) / 336 ** 0.5).to(dev).requires_grad_()
    bias = (torch.randn(128, generator=gen) * 0.1).to(dev).requires_grad_()
    go = torch.randn(B, 21, P_h, P_h, generator=gen).to(dev)

    def pcm_step():
        o = heads.pcm(cam, f, wt, bias)
        o.backward(go)
        cam.grad = wt.grad = bias.grad = None
    ms = timed(pcm_step)
    pairs = B * (P_h * P_h) ** 2
    flops = pairs * 2.0 * (128 + 32) + pairs * 2.0 * (128 + 32) + pairs * 2.0 * (128 + 128 + 128)   # fwd, dcam, bwd
    cam_c, wt_c, b_c, f_c, go_c = cpu_leaf(cam), cpu_leaf(wt), cpu_leaf(bias), f.cpu(), go.cpu()

    def pcm_cpu():
        O.pcm(cam_c, f_c, wt_c, b_c).backward(go_c)
        cam_c.grad = wt_c.grad = b_c.grad = None
    out["pcm_fwd_bwd_config2_stride8"] = beside(dict(ms=ms, pixel_pairs_per_s=pairs / ms * 1e3, executed_tflops=flops / ms / 1e9,
                                                     frac_of_bf16_burst=flops / ms / 1e9 / peaks["bf16_tflops"], bound="tensor"),
                                                pcm_cpu)
    p7 = torch.relu(torch.randn(B, 640, P_h, P_h, generator=gen)).to(dev).requires_grad_()
    fw = (torch.randn(21, 640, generator=gen) * 0.05).to(dev).requires_grad_()
    gcam = torch.randn(B, 21, P_h, P_h, generator=gen).to(dev)

    def cam_step():
        c = heads.cam_head(p7, fw)
        e, l = heads.image_embedding_logits(p7, fw)
        torch.autograd.backward([c, l], [gcam, torch.ones_like(l)])
        p7.grad = fw.grad = None
    ms = timed(cam_step)
    byt = p7.numel() * 4 * 4 + B * 21 * P_h * P_h * 4 * 3      # p7 read x2 (cam, gap), dp7 written x2, cam w + r, dcam r
    p7_c, fw_c, gcam_c = cpu_leaf(p7), cpu_leaf(fw), gcam.cpu()

    def cam_cpu():
        c = O.cam_head(p7_c, fw_c)
        e, l = O.image_embedding_logits(p7_c, fw_c)
        torch.autograd.backward([c, l], [gcam_c, torch.ones_like(l)])
        p7_c.grad = fw_c.grad = None
    out["cam_head_gap_logits_fwd_bwd"] = beside(dict(ms=ms, algorithmic_gb_s=byt / ms / 1e6,
                                                     frac_of_hbm=byt / ms / 1e6 / peaks["hbm_gbs"], bound="hbm"),
                                                cam_cpu)
    # ---- IMC, B=64 images, ch=640
    emb = torch.randn(64, 640, generator=gen).abs().to(dev).requires_grad_()
    lab = workloads.synth_labels(64, gen).to(dev)

    def imc_step():
        l = image_level_contrast(emb, lab)
        if torch.is_tensor(l):
            l.backward()
        emb.grad = None
    emb_c, lab_c = cpu_leaf(emb), lab.cpu()

    def imc_cpu():
        l = O.image_level_contrast(emb_c, lab_c)
        if torch.is_tensor(l):
            l.backward()
        emb_c.grad = None
    out["imc_fwd_bwd_B64"] = beside(dict(ms=timed(imc_step), bound="latency"), imc_cpu)
    # ---- PixPro, B=8, 21x224x224
    fm1 = torch.rand(8, 21, 224, 224, generator=gen).to(dev).requires_grad_()
    fm2 = torch.rand(8, 21, 224, 224, generator=gen).to(dev)
    hw = torch.randint(100, 200, (8, 2), generator=gen)
    c1 = torch.cat([(torch.rand(8, 2, generator=gen) * (224 - hw)).long(), hw], 1)
    c2 = torch.cat([(torch.rand(8, 2, generator=gen) * (224 - hw)).long(), hw], 1)

    def pix_step():
        PixPro(fm1, fm2, c1, c2).backward()
        fm1.grad = None
    ms = timed(pix_step)
    byt = float((hw[:, 0] * hw[:, 1]).sum()) * 21 * 4 * 3 + fm1.numel() * 4
    fm1_c, fm2_c = cpu_leaf(fm1), fm2.cpu()

    def pix_cpu():
        O.pixpro(fm1_c, fm2_c, c1, c2).backward()
        fm1_c.grad = None
    out["pixpro_fwd_bwd_B8"] = beside(dict(ms=ms, algorithmic_gb_s=byt / ms / 1e6, bound="hbm/latency"), pix_cpu)
    # ---- crops + EMD dynamic, B=8
    raw1 = F.avg_pool2d(torch.randn(8, 21, 224, 224, generator=gen), 9, 1, 4).to(dev).requires_grad_()
    raw2 = F.avg_pool2d(torch.randn(8, 21, 224, 224, generator=gen), 9, 1, 4).to(dev)

    def emd_step():
        v1 = F.normalize(heads.cam_softmaxnorm(raw1), dim=1)
        v2 = F.normalize(heads.cam_softmaxnorm(raw2), dim=1)
        np.random.seed(3)
        a, b, _ = get_dynamic_crops(v1, c1, v2.detach(), c2)
        EMD()(a, b, mode="dynamic").backward()
        raw1.grad = None
    raw1_c, raw2_c = cpu_leaf(raw1), raw2.cpu()

    def emd_cpu():
        v1 = F.normalize(O.cam_softmaxnorm(raw1_c), dim=1)
        v2 = F.normalize(O.cam_softmaxnorm(raw2_c), dim=1)
        a, b, _ = O.get_dynamic_crops(v1, c1, v2.detach(), c2, rng=np.random.RandomState(3))
        O.emd_dynamic(a, b).backward()
        raw1_c.grad = None
    out["crops_emd_dynamic_fwd_bwd_B8"] = beside(dict(ms=timed(emd_step), bound="latency"), emd_cpu)
    # ---- static crops + EMD static (the class's default mode), same CAM maps
    from muscle_b200.torchutils import get_static_crops

    def emd_static_step():
        v1 = F.normalize(heads.cam_softmaxnorm(raw1), dim=1)
        v2 = F.normalize(heads.cam_softmaxnorm(raw2), dim=1)
        a, b, _ = get_static_crops(v1, c1, v2.detach(), c2)
        EMD()(a, b).backward()
        raw1.grad = None

    def emd_static_cpu():
        v1 = F.normalize(O.cam_softmaxnorm(raw1_c), dim=1)
        v2 = F.normalize(O.cam_softmaxnorm(raw2_c), dim=1)
        a, b, _ = O.get_static_crops(v1, c1, v2.detach(), c2)
        O.emd_static(a, b)[0].backward()
        raw1_c.grad = None
    out["crops_emd_static_fwd_bwd_B8"] = beside(dict(ms=timed(emd_static_step), bound="latency",
                                                     note="72 window pairs; windows gathered by one launch per view (muscle_crops_fwd), 3 launches for the matching"),
                                                emd_static_cpu)
    # ---- FieldLoss, config 3 scale reduced to B=4 (512x512, 256 ch) to bound memory in this table
    n, h, w, ch = 4, 512, 512, 256
    label = workloads.synth_labels(n, gen)
    label_bg = torch.cat([torch.ones(n, 1), label], 1)
    noise = F.avg_pool2d(torch.randn(n, 21, h, w, generator=gen), 31, 1, 15, count_include_pad=False)
    noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -1e4)
    seg = (noise * 3).to(dev)
    mask = (noise * 2).to(dev)
    dense = torch.randn(n, ch, h, w, generator=gen).to(dev).requires_grad_()
    crit = FieldLoss(sobel_size=5, beta=1e2, k=128)
    lbg = label_bg.to(dev)

    def field_step():
        random.seed(1)
        l, _ = crit(seg, dense, mask, lbg, 7)
        if torch.is_tensor(l):
            l.backward()
        dense.grad = None
    ms = timed(field_step)
    seg_c, mask_c, dense_c = seg.cpu(), mask.cpu(), cpu_leaf(dense)

    def field_cpu():
        l, _ = O.fieldloss_forward(seg_c, dense_c, mask_c, label_bg, step=7, k=128, beta=1e2, sobel_size=5,
                                   rng=random.Random(1))
        if torch.is_tensor(l):
            l.backward()
        dense_c.grad = None
    out["fieldloss_fwd_bwd_B4_512"] = beside(dict(ms=ms, note="includes the one device->host read of the counts and the "
                                                  "memset of the dense gradient (%.0f MB)" % (dense.numel() * 4 / 1e6),
                                                  bound="hbm/latency"), field_cpu, budget_s=10.0, max_reps=2)
    # ---- ER loss at config 2's full resolution: B=8, 21 x 448 x 448, from stride-8 maps
    from muscle_b200.er import er_loss, upsample_bilinear
    cams_s = torch.relu(torch.randn(8, 21, 56, 56, generator=gen)).to(dev)
    sgcs_s = torch.relu(torch.randn(8, 21, 56, 56, generator=gen) + 0.3).to(dev).requires_grad_()
    lab8 = workloads.synth_labels(8, gen).to(dev)

    def er_step():
        l = er_loss(upsample_bilinear(cams_s, (448, 448)), upsample_bilinear(sgcs_s, (448, 448)), lab8)
        l.backward()
        sgcs_s.grad = None
    ms = timed(er_step)
    full = 8 * 21 * 448 * 448 * 4
    byt = 2 * full + 3 * 2 * full + (2 * full + full) + 2 * full      # 2 up-samplings written, 3 select passes, bwd, adjoint
    cams_c, sgcs_c, lab8_c = cams_s.cpu(), cpu_leaf(sgcs_s), lab8.cpu()

    def er_cpu():
        O.er_loss(O.upsample_bilinear(cams_c, (448, 448)), O.upsample_bilinear(sgcs_c, (448, 448)), lab8_c).backward()
        sgcs_c.grad = None
    out["er_loss_upsample_fwd_bwd_B8_448"] = beside(dict(ms=ms, algorithmic_gb_s=byt / ms / 1e6,
                                                         frac_of_hbm=byt / ms / 1e6 / peaks["hbm_gbs"], bound="hbm",
                                                         note="two 8x bilinear up-samplings + three radix-select passes "
                                                              "+ backward + adjoint; no host sync"), er_cpu, max_reps=2)
    # ---- classification losses (8f rank 4), B=16
    from muscle_b200.loss_multilabel import mcl_cls_loss
    lg = (torch.randn(16, 21, generator=gen) * 2).to(dev).requires_grad_()
    lab16 = workloads.synth_labels(16, gen).to(dev)

    def cls_step():
        mcl_cls_loss(lg, lab16)[0].backward()
        lg.grad = None
    lg_c, lab16_c = cpu_leaf(lg), lab16.cpu()

    def cls_cpu():
        O.mcl_cls_loss(lg_c, lab16_c)[0].backward()
        lg_c.grad = None
    out["cls_losses_fwd_bwd_B16"] = beside(dict(ms=timed(cls_step), bound="latency"), cls_cpu)
    if big:
        # ---- FieldLoss at BASELINE config 3's size: B=16, 512x512, 256 channels, k=128, step=7 — the materialised
        # path (the trainer's: dense_ft [16,256,512,512] exists) and the decoder-tail path (8f rank 2: from p3_dec)
        n, hs, ws, ch, h, w = 16, 64, 64, 256, 512, 512
        label = workloads.synth_labels(n, gen)
        label_bg = torch.cat([torch.ones(n, 1), label], 1)
        noise = F.avg_pool2d(torch.randn(n, 21, hs, ws, generator=gen), 5, 1, 2, count_include_pad=False)
        noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -3.0)
        wmat = (torch.randn(21, ch, generator=gen) / ch ** 0.5)
        bias2 = 0.1 * torch.randn(21, generator=gen)
        p3 = (torch.einsum("ck,nkhw->nchw", torch.linalg.pinv(wmat), noise - bias2[None, :, None, None])
              + 0.2 * torch.randn(n, ch, hs, ws, generator=gen)).to(dev).requires_grad_()
        mask16 = (F.interpolate(noise, size=(h, w), mode="bilinear", align_corners=True) * 0.7).to(dev)
        wd, bd, lbg16 = wmat.to(dev), bias2.to(dev), label_bg.to(dev)
        crit16 = FieldLoss(sobel_size=5, beta=1e2, k=128)

        def tail_step():
            random.seed(1)
            l, _, _ = crit16.forward_from_decoder(p3, wd, bd, mask16, lbg16, 7, (h, w))
            if torch.is_tensor(l):
                l.backward()
            p3.grad = None
        ms_tail = timed(tail_step)

        def mat_step():     # what the reference's data flow costs on this path: dense_ft and seg_map at full resolution
            random.seed(1)
            dense16 = upsample_bilinear(p3, (h, w))
            seg16 = torch.nn.functional.conv2d(dense16.detach(), wd[:, :, None, None], bd)
            l, _ = crit16(seg16, dense16, mask16, lbg16, 7)
            if torch.is_tensor(l):
                l.backward()
            p3.grad = None
        ms_mat = timed(mat_step)
        out["fieldloss_decoder_tail_fwd_bwd_B16_512"] = dict(
            ms=ms_tail, bound="hbm/latency",
            note="config 3: p3_dec [16,256,64,64] -> loss + d p3_dec; dense_ft [16,256,512,512] (4.3 GB) never formed")
        out["fieldloss_materialised_fwd_bwd_B16_512"] = dict(
            ms=ms_mat, bound="hbm",
            note="same inputs with dense_ft up-sampled to full resolution first (the reference's data flow; fuse_dec by "
                 "torch conv2d), then the ordinary FieldLoss and the adjoint of the up-sampling")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    print(json.dumps(collect(args.iters, not args.no_cpu)))


if __name__ == "__main__":
    main()
