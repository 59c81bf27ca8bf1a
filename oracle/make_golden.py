"""Mint golden fixtures by running the UNMODIFIED reference (authoring container only).

TEST INFRASTRUCTURE ONLY.  Usage:  python -m oracle.make_golden
Writes tests/golden/*.pt — seeded inputs plus the outputs (and input-gradients) that
the reference's own functions return for them on CPU.  tests/test_oracle_golden.py
checks oracle/muscle_oracle.py against these; the GPU parity tests check the CUDA path
against the oracle and against these same files.
"""
import os
import random
import sys

import numpy as np
import torch
import torch.nn.functional as F

from oracle import reference_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _blob_classes(b, h, w, n_cls, gen, smooth=7):
    noise = torch.randn(b, n_cls, h, w, generator=gen)
    noise = F.avg_pool2d(noise, smooth, 1, smooth // 2, count_include_pad=False)
    return noise.argmax(1)


def gen_info_nce(src):
    out = {}
    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        g = torch.Generator().manual_seed(11)
        q = F.normalize(torch.randn(64, 32, generator=g, dtype=dt), dim=1).requires_grad_()
        p = F.normalize(torch.randn(16, 32, generator=g, dtype=dt), dim=1).requires_grad_()
        k = F.normalize(torch.randn(96, 32, generator=g, dtype=dt), dim=1).requires_grad_()
        loss = src.loss_multilabel.info_nce(q, p, k)
        gq, gp, gk = torch.autograd.grad(loss, (q, p, k))
        none = src.loss_multilabel.info_nce(q, p, k, temperature=0.07, reduction="none")
        out[tag] = dict(q=q.detach(), p=p.detach(), k=k.detach(), loss=loss.detach(), gq=gq, gp=gp, gk=gk,
                        none_t007=none.detach())
    return out


def gen_info_nce_anynorm(src):
    """The reference primitive on rows that are NOT unit-norm (it never normalises, src/loss_multilabel.py:9-21), with
    reduction='none' and an upstream gradient per row."""
    out = {}
    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        g = torch.Generator().manual_seed(29)
        d = 48
        q = (torch.randn(200, d, generator=g, dtype=dt) * (0.3 + torch.rand(200, 1, generator=g, dtype=dt)) / d ** 0.5).requires_grad_()
        p = ((q[:1].detach() + 0.4 * torch.randn(9, d, generator=g, dtype=dt) / d ** 0.5) * 0.8).requires_grad_()
        k = (torch.randn(333, d, generator=g, dtype=dt) * (0.2 + 1.1 * torch.rand(333, 1, generator=g, dtype=dt)) / d ** 0.5).requires_grad_()
        up = torch.randn(200, generator=g, dtype=dt)
        loss = src.loss_multilabel.info_nce(q, p, k, temperature=0.3)
        gq, gp, gk = torch.autograd.grad(loss, (q, p, k))
        rows = src.loss_multilabel.info_nce(q, p, k, temperature=0.3, reduction="none")
        nq, np_, nk = torch.autograd.grad((rows * up).sum(), (q, p, k))
        out[tag] = dict(q=q.detach(), p=p.detach(), k=k.detach(), up=up, loss=loss.detach(), gq=gq, gp=gp, gk=gk,
                        rows=rows.detach(), nq=nq, np=np_, nk=nk)
    return out


def gen_dense_infonce(src):
    """The composition of SURVEY §8c evaluated with the reference primitive only."""
    out = {}
    for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        g = torch.Generator().manual_seed(23)
        b, d, h, w, c = 2, 32, 12, 12, 5
        cls = _blob_classes(b, h, w, c, g, smooth=5).reshape(-1)
        cls[torch.rand(cls.shape, generator=g) < 0.05] = -1          # ignored pixels
        x = torch.randn(b * h * w, d, generator=g, dtype=dt)
        x = (x + 0.7 * torch.randn(c, d, generator=g, dtype=dt)[cls.clamp_min(0)]).requires_grad_()
        e = F.normalize(x, dim=1)
        valid = cls >= 0
        per_query = []
        for ci in sorted(set(cls[valid].tolist())):
            sel = cls == ci
            neg = valid & ~sel
            per_query.append(src.loss_multilabel.info_nce(e[sel], e[sel], e[neg], temperature=0.1,
                                                          reduction="none"))
        loss = torch.cat(per_query).mean()
        gx, = torch.autograd.grad(loss, x)
        out[tag] = dict(x=x.detach(), cls=cls, loss=loss.detach(), gx=gx)
    # the same composition at the projection dims the tensor-pipe passes take (so that the CUDA path itself,
    # not only the oracle, is checked against the reference primitive): x in fp32, reference run in fp32 and fp64
    for d in (64, 128):
        g = torch.Generator().manual_seed(23 + d)
        b, h, w, c = 3, 14, 15, 6
        cls = _blob_classes(b, h, w, c, g, smooth=5).reshape(-1)
        cls[torch.rand(cls.shape, generator=g) < 0.04] = -1
        x32 = torch.randn(b * h * w, d, generator=g)
        x32 = x32 + 0.7 * torch.randn(c, d, generator=g)[cls.clamp_min(0)]
        rec = dict(x=x32, cls=cls, num_classes=c)
        for dt, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
            x = x32.to(dt).clone().requires_grad_()
            e = F.normalize(x, dim=1)
            valid = cls >= 0
            per_query = []
            for ci in sorted(set(cls[valid].tolist())):
                sel = cls == ci
                per_query.append(src.loss_multilabel.info_nce(e[sel], e[sel], e[valid & ~sel], temperature=0.1,
                                                              reduction="none"))
            loss = torch.cat(per_query).mean()
            gx, = torch.autograd.grad(loss, x)
            rec["loss_" + tag] = loss.detach()
            rec["gx_" + tag] = gx
        out["d%d" % d] = rec
    return out


def gen_imc(src):
    g = torch.Generator().manual_seed(5)
    emb = torch.randn(8, 48, generator=g).abs().requires_grad_()
    label = torch.zeros(8, 20)
    sets = [[14], [14], [3], [3, 7], [14], [6], [14, 6], [2]]
    for i, s in enumerate(sets):
        label[i, s] = 1
    loss = src.loss_multilabel.image_level_contrast(emb, label)
    ge, = torch.autograd.grad(loss, emb)
    # a batch where nothing qualifies -> python 0
    label0 = torch.zeros(3, 20)
    label0[:, 1] = 1
    zero = src.loss_multilabel.image_level_contrast(torch.randn(3, 48, generator=g), label0)
    g2 = torch.Generator().manual_seed(6)
    emb2 = torch.randn(16, 384, generator=g2).abs().requires_grad_()
    label2 = (torch.rand(16, 20, generator=g2) < 0.08).float()
    label2[:, 14] = (torch.rand(16, generator=g2) < 0.4).float()
    loss2 = src.loss_multilabel.image_level_contrast(emb2, label2)
    ge2, = torch.autograd.grad(loss2, emb2) if torch.is_tensor(loss2) else (torch.zeros_like(emb2),)
    return dict(emb=emb.detach(), label=label, loss=loss.detach(), ge=ge, zero_is_tensor=torch.is_tensor(zero),
                zero_value=float(zero), emb2=emb2.detach(), label2=label2,
                loss2=loss2.detach() if torch.is_tensor(loss2) else None, ge2=ge2)


def gen_pixpro(src):
    g = torch.Generator().manual_seed(7)
    fm1 = torch.rand(3, 21, 28, 28, generator=g).requires_grad_()
    fm2 = torch.rand(3, 21, 28, 28, generator=g)
    fm1.data[:, 5] = 0  # an all-zero channel column set exercises the eps clamp
    c1 = torch.tensor([[0, 4, 20, 18], [6, 0, 15, 28], [3, 3, 7, 9]])
    c2 = torch.tensor([[8, 0, 20, 18], [0, 0, 15, 28], [20, 19, 7, 9]])
    loss = src.loss_multilabel.PixPro(fm1, fm2, c1, c2)
    g1, = torch.autograd.grad(loss, fm1)
    return dict(fm1=fm1.detach(), fm2=fm2, c1=c1, c2=c2, loss=loss.detach(), g1=g1)


def gen_emd(src):
    g = torch.Generator().manual_seed(9)
    raw1 = torch.randn(3, 21, 56, 56, generator=g)
    raw2 = torch.randn(3, 21, 56, 56, generator=g)
    raw1 = F.avg_pool2d(raw1, 5, 1, 2).requires_grad_()
    raw2 = F.avg_pool2d(raw2, 5, 1, 2)
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(raw2 * 4, 1), dim=1)
    c1 = torch.tensor([[0, 8, 48, 44], [10, 0, 33, 56], [0, 0, 10, 40]])
    c2 = torch.tensor([[8, 0, 48, 44], [0, 0, 33, 56], [30, 5, 10, 40]])
    np.random.seed(1234)
    crops1, crops2, bidx = src.torchutils.get_dynamic_crops(vw1, c1, vw2.detach(), c2)
    loss = src.loss_multilabel.EMD()(crops1, crops2, mode="dynamic")
    g1, = torch.autograd.grad(loss, raw1)
    # one isolated Sinkhorn call for the kernel-level test
    emd = src.loss_multilabel.EMD()
    x = crops1[0][0].reshape(1, 21, -1).permute(2, 1, 0)
    y = crops2[0][1].reshape(1, 21, -1).permute(2, 1, 0)
    cost = emd.pair_wise_cos(x, y).squeeze(0).detach()
    w1 = emd.get_weight_vector(crops1[0][0], crops2[0][1]).detach()
    w2 = emd.get_weight_vector(crops2[0][1], crops1[0][0]).detach()
    sk = emd.sinkhorn_logsumexp(cost, w1, w2, maxiter=10, grad=False)
    return dict(raw1=raw1.detach(), raw2=raw2, c1=c1, c2=c2, np_seed=1234, batch_indices=bidx,
                n_crops1=[len(c) for c in crops1], n_crops2=[len(c) for c in crops2],
                crop_shapes1=[[tuple(t.shape) for t in c] for c in crops1],
                crop_shapes2=[[tuple(t.shape) for t in c] for c in crops2],
                loss=loss.detach(), g1=g1, sk_cost=cost, sk_w1=w1, sk_w2=w2, sk_value=sk)


def gen_emd_static(src):
    """get_static_crops + EMD()(…, mode='static') (src/torchutils.py:177-214, src/loss_multilabel.py:259-285)."""
    g = torch.Generator().manual_seed(15)
    raw1 = F.avg_pool2d(torch.randn(3, 21, 40, 40, generator=g), 5, 1, 2).requires_grad_()
    raw2 = F.avg_pool2d(torch.randn(3, 21, 40, 40, generator=g), 5, 1, 2)
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(raw2 * 4, 1), dim=1)
    c1 = torch.tensor([[0, 6, 34, 22], [4, 0, 17, 40], [0, 0, 10, 30]])      # third box is skipped (h < 15)
    c2 = torch.tensor([[6, 0, 34, 22], [0, 0, 17, 40], [20, 5, 10, 30]])
    crops1, crops2, bidx = src.torchutils.get_static_crops(vw1, c1, vw2.detach(), c2)
    emd = src.loss_multilabel.EMD()
    loss = emd(crops1, crops2)                                                # default mode is 'static'
    g1, = torch.autograd.grad(loss, raw1)
    dists = []
    for a, b in zip(crops1, crops2):
        dists.append(emd.static_matching([a, a], [b, b]).detach() / 2)
    return dict(raw1=raw1.detach(), raw2=raw2, c1=c1, c2=c2, batch_indices=bidx,
                crop_shapes1=[tuple(t.shape) for t in crops1], crop_shapes2=[tuple(t.shape) for t in crops2],
                loss=loss.detach(), g1=g1, dists=torch.stack(dists))


def gen_heads(src):
    """CAM head + embedding + PCM evaluated in situ by the reference model (b3, random weights)."""
    torch.manual_seed(3)
    model = src.MuSCLe(num_classes=21, pretrained="efficientnet-b3", last_pooling=False, mode="enc")
    model.eval()
    x = torch.randn(2, 3, 64, 64)
    with torch.no_grad():
        fps = model.backbone(x)
    p1, p3, p5, p7 = (fps[i].detach() for i in (model.p1_seq, model.p3_seq, model.p5_seq, model.p7_seq))
    cams, sgc, emb, logits = model(x, cam="cam")
    # PCM alone on leaf inputs so that gradients can be compared
    g = torch.Generator().manual_seed(4)
    n, h, w = 2, 10, 12
    cin = model.fuse.weight.shape[1]
    cam_in = torch.rand(n, 21, h, w, generator=g).requires_grad_()
    f_in = torch.relu(torch.randn(n, cin, h, w, generator=g))
    model.zero_grad()
    out = model.PCM(cam_in, f_in)
    go = torch.randn(out.shape, generator=g)
    (out * go).sum().backward()
    return dict(p1=p1, p3=p3, p5=p5, p7=p7, fc_w=model.fc.weight.detach().clone(),
                fuse_w=model.fuse.weight.detach().clone(), fuse_b=model.fuse.bias.detach().clone(),
                cams_up=cams.detach(), sgc_up=sgc.detach(), emb=emb.detach(), logits=logits.detach(),
                img_hw=(64, 64),
                pcm_cam=cam_in.detach(), pcm_f=f_in, pcm_out=out.detach(), pcm_go=go,
                pcm_gcam=cam_in.grad.clone(), pcm_gw=model.fuse.weight.grad.clone(),
                pcm_gb=model.fuse.bias.grad.clone())


def gen_fieldloss(src):
    g = torch.Generator().manual_seed(13)
    n, c, h, w, ch, k, step = 2, 21, 44, 44, 24, 12, 3
    label = torch.zeros(n, 20)
    label[0, [2, 14]] = 1
    label[1, [14]] = 1
    label_bg = torch.cat([torch.ones(n, 1), label], 1)
    noise = F.avg_pool2d(torch.randn(n, c, h, w, generator=g), 9, 1, 4, count_include_pad=False)
    noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -1e4)
    seg = (noise * 3 + 0.05 * torch.randn(n, c, h, w, generator=g)).requires_grad_()
    mask = noise * 2 + 0.3 * torch.randn(n, c, h, w, generator=g)
    dense = torch.randn(n, ch, h, w, generator=g).requires_grad_()
    crit = src.edge.FieldLoss(sobel_size=5, beta=1e2, k=k)
    random.seed(77)
    loss, mag = crit(seg, dense, mask, label_bg, step)
    assert torch.is_tensor(loss), "fixture must exercise the bifilter"
    gd, = torch.autograd.grad(loss, dense)
    # stage outputs
    with torch.no_grad():
        edges = crit.fg_edge(seg, label_bg, reduction=False)
        mag2, orient = crit.quantize(edges.clone())
        mx = torch.max(mag2.view(n, 20, -1), dim=-1)[0].unsqueeze(2).unsqueeze(2)
        pos = (mag2 >= 0.8 * mx) & (mx > 1)
        pos = pos * label_bg[:, 1:].unsqueeze(2).unsqueeze(2).bool()
        out_i, in_i = crit.in_out_div(orient[0, 14] + 1, pos[0, 14], step_outs=[step], step_ins=[step])
    # gate cases: no fg class present -> False (edge.py:377-378); enough boundary pixels but
    # in/out lists never longer than k -> python 0 (edge.py:359-361)
    crit2 = src.edge.FieldLoss(sobel_size=5, beta=1e2, k=k)
    flat = torch.zeros(1, c, 24, 24)
    only_bg = torch.cat([torch.ones(1, 1), torch.zeros(1, 20)], 1)
    loss0, _ = crit2(flat, torch.randn(1, ch, 24, 24, generator=g), flat, only_bg, step)
    crit3 = src.edge.FieldLoss(sobel_size=5, beta=1e2, k=100000)
    loss1, _ = crit3(seg.detach(), dense.detach(), mask, label_bg, step)
    return dict(seg=seg.detach(), dense=dense.detach(), mask=mask, label_bg=label_bg, k=k, step=step,
                beta=1e2, py_seed=77, loss=loss.detach(), mag=mag, gd=gd, edges=edges, orient=orient,
                pos=pos, out_idx_0_14=out_i, in_idx_0_14=in_i, gate_false=(loss0 is False), big_k_is_zero=(not torch.is_tensor(loss1) and loss1 == 0))


def gen_fieldloss_sobel3(src):
    """FieldLoss built with sobel_size=3 (src/edge.py:33-35; the trainer passes 5, the constructor accepts both)."""
    g = torch.Generator().manual_seed(31)
    n, c, h, w, ch, k, step = 2, 21, 40, 48, 16, 10, 2
    label = torch.zeros(n, 20)
    label[0, [0, 7]] = 1
    label[1, [7, 11]] = 1
    label_bg = torch.cat([torch.ones(n, 1), label], 1)
    noise = F.avg_pool2d(torch.randn(n, c, h, w, generator=g), 9, 1, 4, count_include_pad=False)
    noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -1e4)
    seg = noise * 3 + 0.05 * torch.randn(n, c, h, w, generator=g)
    mask = noise * 2 + 0.3 * torch.randn(n, c, h, w, generator=g)
    dense = torch.randn(n, ch, h, w, generator=g).requires_grad_()
    crit = src.edge.FieldLoss(sobel_size=3, beta=1e2, k=k)
    random.seed(91)
    loss, mag = crit(seg, dense, mask, label_bg, step)
    assert torch.is_tensor(loss), "fixture must exercise the bifilter"
    gd, = torch.autograd.grad(loss, dense)
    with torch.no_grad():
        edges = crit.fg_edge(seg, label_bg, reduction=False)
        _, orient = crit.quantize(edges.clone())
    return dict(seg=seg, dense=dense.detach(), mask=mask, label_bg=label_bg, k=k, step=step, beta=1e2, py_seed=91,
                loss=loss.detach(), mag=mag, gd=gd, orient=orient.to(torch.uint8))


def _reference_functions(rel_path, names):
    """Compiles the named top-level functions of a reference SCRIPT (which cannot be imported: it pulls
    in data / cv2 / tensorboardX at module level) from where it lies, and returns them."""
    import ast
    path = os.path.join(reference_loader.REFERENCE_ROOT, rel_path)
    with open(path) as f:
        text = f.read()
    tree = ast.parse(text)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(keep) == len(names), (rel_path, names)
    mod = ast.Module(body=keep, type_ignores=[])
    ns = {"torch": torch, "F": F, "np": np}
    exec(compile(mod, path, "exec"), ns)
    return [ns[n] for n in names]


def _reference_lines(rel_path, first, last):
    """Source lines [first, last] (1-based) of a reference file, dedented, for exec in a prepared namespace."""
    import textwrap
    path = os.path.join(reference_loader.REFERENCE_ROOT, rel_path)
    with open(path) as f:
        lines = f.readlines()
    return compile(textwrap.dedent("".join(lines[first - 1:last])), path + ":%d-%d" % (first, last), "exec")


def gen_cam_norms(src):
    """train_mcl.py:21-36 (cam_maxnorm, cam_softmaxnorm) and the model's own variant src/MuSCLe.py:202-211."""
    cam_maxnorm, cam_softmaxnorm = _reference_functions("train_mcl.py", ["cam_maxnorm", "cam_softmaxnorm"])
    g = torch.Generator().manual_seed(31)
    x = torch.randn(3, 21, 13, 11, generator=g)
    x[0, 4] = 0.0                      # an all-zero plane: min == max
    x[1, 7] = -x[1, 7].abs()           # an all-negative plane: relu -> all zero
    out = {}
    for name, fn in (("maxnorm", cam_maxnorm), ("softmaxnorm", cam_softmaxnorm)):
        xi = x.clone().requires_grad_()
        y = fn(xi)
        go = torch.randn(y.shape, generator=g)
        gx, = torch.autograd.grad((y * go).sum(), xi)
        out[name] = dict(y=y.detach(), go=go, gx=gx)
    out["x"] = x
    return out


def gen_er_loss(src):
    """The ER loss of train_mcl.py:175-188 with the up-sampling of src/MuSCLe.py:256-257 in front of it: the
    reference's own source lines are executed on seeded stand-ins for the model outputs."""
    cam_softmaxnorm, = _reference_functions("train_mcl.py", ["cam_softmaxnorm"])
    g = torch.Generator().manual_seed(37)
    n, h, w, H, W = 3, 7, 9, 28, 36
    cams_small = torch.relu(torch.randn(n, 21, h, w, generator=g))
    SGC_small = torch.relu(torch.randn(n, 21, h, w, generator=g) + 0.3).requires_grad_()
    label = torch.zeros(n, 20)
    label[0, [1, 14]] = 1
    label[1, [14]] = 1
    label[2, [3, 7, 19]] = 1
    ns = dict(torch=torch, F=F, cams=cams_small, SGC=SGC_small, H=H, W=W, label=label, cam_softmaxnorm=cam_softmaxnorm)
    exec(_reference_lines("src/MuSCLe.py", 256, 257), ns)            # cams, SGC up-sampled (align_corners=True)
    ns["raw_cams"], ns["raw_sgcs"] = ns["cams"], ns["SGC"]
    up_cams, up_sgc = ns["cams"].detach().clone(), ns["SGC"].detach().clone()
    exec(_reference_lines("train_mcl.py", 167, 170), ns)             # label_with_bg
    exec(_reference_lines("train_mcl.py", 175, 178), ns)             # cams / sgcs normalised, valid_channel
    exec(_reference_lines("train_mcl.py", 185, 188), ns)             # label mask, top-k mean
    loss_er = ns["loss_er"]
    g_small, = torch.autograd.grad(loss_er, SGC_small)
    return dict(cams_small=cams_small, sgc_small=SGC_small.detach(), label=label, size=(H, W), up_cams=up_cams,
                up_sgc=up_sgc, valid_channel=ns["valid_channel"], k=int(0.2 * ns["valid_channel"] * H * W),
                loss_er=loss_er.detach(), g_sgc_small=g_small)


def gen_cls_losses(src):
    """FocalLoss, Log_Sum_Exp_Pairwise_Loss (src/loss_multilabel.py:24-33,68-91) and the soft-margin term, composed
    as train_mcl.py:180-183 composes them."""
    g = torch.Generator().manual_seed(41)
    logits = (torch.randn(6, 21, generator=g) * 2).requires_grad_()
    label = (torch.rand(6, 20, generator=g) < 0.15).float()
    label[0] = 0
    label[0, 14] = 1
    import torch.nn as nn
    ns = dict(torch=torch, logits=logits, label=label, criterion1=src.loss_multilabel.FocalLoss(),
              criterion2=src.loss_multilabel.Log_Sum_Exp_Pairwise_Loss, criterion3=nn.MultiLabelSoftMarginLoss())
    exec(_reference_lines("train_mcl.py", 180, 183), ns)
    out = dict(logits=logits.detach(), label=label)
    for name in ("loss_focal", "loss_softmargin", "loss_pair", "loss_cls"):
        out[name] = ns[name].detach()
    out["g_logits"], = torch.autograd.grad(ns["loss_cls"], logits)
    out["pair_rows"] = src.loss_multilabel.Log_Sum_Exp_Pairwise_Loss(torch.sigmoid(logits[:, 1:]), label).detach()
    return out


def gen_decoder_tail(src):
    """The decoder tail of src/MuSCLe.py:285-286 (8x bilinear up-sampling of p3_dec, align_corners=True, then the
    1x1 fuse_dec) feeding FieldLoss the way train_muscle.py:196 does: loss and gradients w.r.t. p3_dec and
    fuse_dec.  p3_dec is built so that the segmentation has blob structure (boundaries for the loss to sample)."""
    import torch.nn as nn
    g = torch.Generator().manual_seed(43)
    n, ch, hs, ws, up, k, step = 2, 32, 11, 12, 4, 10, 3
    H, W = hs * up, ws * up
    torch.manual_seed(43)
    fuse_dec = nn.Conv2d(ch, 21, 1)
    label = torch.zeros(n, 20)
    label[0, [2, 14]] = 1
    label[1, [14]] = 1
    label_bg = torch.cat([torch.ones(n, 1), label], 1)
    noise = F.avg_pool2d(torch.randn(n, 21, hs, ws, generator=g), 5, 1, 2, count_include_pad=False)
    noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -3.0)     # seg_map in [-3, 1]: beta = 100 sharpens it,
    wmat = fuse_dec.weight.detach().reshape(21, ch)                          # and p3_dec stays O(1) (live channel softmax)
    target = noise - fuse_dec.bias.detach()[None, :, None, None]
    p3 = torch.einsum("ck,nkhw->nchw", torch.linalg.pinv(wmat), target) + 0.2 * torch.randn(n, ch, hs, ws, generator=g)
    p3_dec = p3.clone().requires_grad_()
    mask = F.interpolate(noise, size=(H, W), mode="bilinear", align_corners=True) * 0.7 + 0.3 * torch.randn(n, 21, H, W, generator=g)

    class _Self:            # the two source lines use self.fuse_dec only
        pass
    me = _Self()
    me.fuse_dec = fuse_dec
    ns = dict(F=F, self=me, p3_dec=p3_dec, H=H, W=W)
    exec(_reference_lines("src/MuSCLe.py", 285, 286), ns)
    seg_map, dense_ft = ns["seg_map"], ns["dense_ft"]
    crit = src.edge.FieldLoss(sobel_size=5, beta=1e2, k=k)
    random.seed(91)
    loss, mag = crit(seg_map, dense_ft, mask, label_bg, step)
    assert torch.is_tensor(loss), "fixture must exercise the bifilter"
    gp, gw, gb = torch.autograd.grad(loss, (p3_dec, fuse_dec.weight, fuse_dec.bias), allow_unused=True)
    return dict(p3_dec=p3_dec.detach(), fuse_w=fuse_dec.weight.detach().clone(), fuse_b=fuse_dec.bias.detach().clone(),
                mask=mask, label_bg=label_bg, size=(H, W), k=k, step=step, beta=1e2, py_seed=91,
                seg_map=seg_map.detach(), dense_ft=dense_ft.detach(), loss=loss.detach(), mag=mag,
                g_p3=gp, g_w=None if gw is None else gw, g_b=None if gb is None else gb)


def gen_irn(src):
    """IRN random walk: src/indexing.py:116-150 (to_transition_matrix, propagate_to_edge; the latter hard-codes
    .cuda(), made the identity for this CPU run)."""
    import src.indexing as indexing
    g = torch.Generator().manual_seed(47)
    hw = 96
    a = torch.rand(hw, hw, generator=g)
    aff = 0.5 * (a + a.t())
    aff.fill_diagonal_(1.0)
    out = dict(aff=aff)
    for times in (1, 4, 8):
        out["trans_%d" % times] = indexing.to_transition_matrix(aff, beta=10, times=times)
    out["trans_b2_3"] = indexing.to_transition_matrix(aff, beta=2, times=3)
    h, w = 12, 14
    x = torch.rand(3, h, w, generator=g)
    edge = torch.sigmoid(4 * F.avg_pool2d(torch.randn(1, 1, h, w, generator=g), 3, 1, 1))[0, 0]
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        rw = indexing.propagate_to_edge(x.clone(), edge.clone(), radius=5, beta=10, exp_times=8)
    finally:
        torch.Tensor.cuda = orig
    out.update(x=x, edge=edge, rw=rw)
    return out


def main():
    src = reference_loader.load()
    os.makedirs(OUT, exist_ok=True)
    jobs = dict(info_nce=gen_info_nce, info_nce_anynorm=gen_info_nce_anynorm, dense_infonce=gen_dense_infonce, imc=gen_imc, pixpro=gen_pixpro,
                emd=gen_emd, emd_static=gen_emd_static, heads=gen_heads, fieldloss=gen_fieldloss,
                fieldloss_sobel3=gen_fieldloss_sobel3, cam_norms=gen_cam_norms,
                er_loss=gen_er_loss, cls_losses=gen_cls_losses, decoder_tail=gen_decoder_tail, irn=gen_irn)
    only = sys.argv[1:]
    for name, fn in jobs.items():
        if only and name not in only:
            continue
        torch.manual_seed(0)
        data = fn(src)
        path = os.path.join(OUT, name + ".pt")
        torch.save(data, path)
        print("%-14s -> %s (%.1f KiB)" % (name, path, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
