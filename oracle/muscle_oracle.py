"""CPU oracle for the MuSCLe contrastive-loss hot path (TEST INFRASTRUCTURE ONLY).

A restatement, in plain torch-on-CPU tensor algebra, of what the reference
computes on the path SURVEY.md §8(a) lists.  Every function cites the reference
file:line it follows (paths relative to the upstream repo root).  It is never
imported by the product package ``muscle_b200`` — only by ``tests/``,
``__graft_entry__.smoke()`` and the CPU-baseline legs of ``bench.py``.

Parity status: **pinned against the reference itself**.  The reference has no
tests or golden vectors of its own, so ``oracle/make_golden.py``
imports the unmodified reference from ``/root/reference`` (authoring container
only), runs its functions on seeded inputs and stores input/output pairs under
``tests/golden/``; ``tests/test_oracle_golden.py`` checks this file against
those fixtures.  The one composition the reference leaves open — the dense
masked InfoNCE over a whole batch, built from the reference's ``info_nce``
primitive — is pinned the same way (the fixture is produced by calling the
reference ``info_nce`` once per class, see ``make_golden.py``).

All functions are dtype-generic (run them in float64 for "truth", float32 for
the reference's own fp32 path) and differentiable through torch autograd, which
is exactly how the reference obtains its gradients.
"""
from __future__ import annotations

import math
import random as _pyrandom

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------
# a1  InfoNCE primitive and the dense, label-masked composition
# --------------------------------------------------------------------------


def info_nce(query, positive_keys, negative_keys, temperature=0.1, reduction="mean"):
    """src/loss_multilabel.py:9-21.

    positive logit  = query . mean_over_rows(positive_keys)      (:12)
    negative logits = query @ negative_keys^T                    (:15)
    loss_q = logsumexp([pos, negs] / T) - pos / T                (:18-21, CE with target 0)
    Inputs are NOT normalised here (the caller does that).
    """
    proto = positive_keys.mean(dim=0)
    pos = (query @ proto) / temperature
    neg = (query @ negative_keys.t()) / temperature
    lse = torch.logsumexp(torch.cat([pos[:, None], neg], dim=1), dim=1)
    per_query = lse - pos
    if reduction == "none":
        return per_query
    if reduction == "sum":
        return per_query.sum()
    return per_query.mean()


def l2_normalize_pixels(x, eps=1e-12):
    """Row-wise x / max(||x||, eps) — F.normalize semantics, used by the callers of
    info_nce-style losses (train_mcl.py:218-219 normalises its views this way)."""
    return x / x.norm(dim=-1, keepdim=True).clamp_min(eps)


def class_prototypes(ehat, cls, num_classes, weights=None):
    """Mean (or weighted mean) of pixel embeddings per class.

    Hard variant == ``positive_keys.mean`` of src/loss_multilabel.py:12 with
    positive_keys = rows of class c.  Soft variant: sum_p w[p,c] e_p / sum_p w[p,c]
    (SURVEY.md §8c); with a 0/1 weight it reduces to the hard one.
    Returns (proto [C,D], mass [C])."""
    if weights is None:
        valid = cls >= 0
        onehot = torch.zeros(ehat.shape[0], num_classes, dtype=ehat.dtype)
        onehot[valid, cls[valid]] = 1
        weights = onehot
    mass = weights.sum(0)
    proto = (weights.t() @ ehat) / mass.clamp_min(1e-30)[:, None]
    return proto, mass


def dense_infonce(ehat_q, cls_q, ehat_k=None, cls_k=None, temperature=0.1,
                  num_classes=None, proto=None, n_valid=None, row_chunk=4096):
    """Dense label-masked pixel InfoNCE (composition fixed in SURVEY.md §8c).

    For every valid query pixel q of class c:
        pos_q  = e_q . proto_c                       (pixel -> class prototype)
        negs_q = { e_q . e_k : cls_k != c, k valid } (pixel -> pixel)
        loss_q = info_nce(...) of src/loss_multilabel.py:9-21
    and the result is the mean over all valid queries.  Equivalent to calling the
    reference ``info_nce(E[cls==c], E[cls==c], E[cls!=c], T, 'none')`` per class and
    averaging the concatenation.  Keys default to the queries (self-similarity);
    pass the gathered global keys for the multi-GPU form.  ``cls < 0`` marks pixels
    that are neither queries nor keys.  Row-chunked so [Q,K] never has to fit.
    """
    if ehat_k is None:
        ehat_k, cls_k = ehat_q, cls_q
    if num_classes is None:
        num_classes = int(max(cls_q.max(), cls_k.max())) + 1
    if proto is None:
        proto, _ = class_prototypes(ehat_k, cls_k, num_classes)
    vq = cls_q >= 0
    vk = cls_k >= 0
    total = ehat_q.new_zeros(())
    for s in range(0, ehat_q.shape[0], row_chunk):
        q = ehat_q[s:s + row_chunk]
        cq = cls_q[s:s + row_chunk]
        ok = vq[s:s + row_chunk]
        pos = (q * proto[cq.clamp_min(0)]).sum(-1) / temperature
        logits = (q @ ehat_k.t()) / temperature
        neg_mask = (cq[:, None] != cls_k[None, :]) & vk[None, :]
        logits = logits.masked_fill(~neg_mask, float("-inf"))
        lse = torch.logsumexp(torch.cat([pos[:, None], logits], dim=1), dim=1)
        total = total + ((lse - pos) * ok.to(q.dtype)).sum()
    if n_valid is None:
        n_valid = vq.sum()
    return total / n_valid


def dense_infonce_value_and_grad_chunked(x, cls, temperature=0.1, num_classes=None, row_chunk=2048):
    """(loss, d loss / d x) of ``dense_infonce_from_features`` at sizes whose [Q,K] logits do not fit: the loss is
    a sum over query chunks, so each chunk's share is back-propagated on its own (through the normalisation and the
    prototypes, like the whole) and the gradients are added — autograd exactly as the reference would obtain
    them, only never holding more than one [row_chunk, K] block."""
    x = x.detach().clone().requires_grad_()
    n = x.shape[0]
    n_valid = (cls >= 0).sum()
    total = 0.0
    for s0 in range(0, n, row_chunk):
        e = l2_normalize_pixels(x)
        if num_classes is None:
            num_classes = int(cls.max()) + 1
        proto, _ = class_prototypes(e, cls, num_classes)
        part = dense_infonce(e[s0:s0 + row_chunk], cls[s0:s0 + row_chunk], e, cls, temperature, num_classes,
                             proto=proto, n_valid=n_valid, row_chunk=row_chunk)
        part.backward()
        total += float(part.detach())
    return total, x.grad.detach()


def dense_infonce_per_class(ehat, cls, temperature=0.1, query_idx=None):
    """The composition of SURVEY.md §8c written exactly the way the reference primitive would be
    called: ONE ``info_nce(E[cls==c], E[cls==c], E[cls!=c], T, 'none')`` per class present
    (src/loss_multilabel.py:9-21), i.e. only cross-class pairs are ever multiplied; the per-query
    losses are concatenated and averaged over the valid queries.  ``query_idx`` restricts the
    QUERIES to a sample (keys and prototypes stay the whole batch): the result is then the sum of the
    sampled queries' losses / n_valid — the bounded-sample form bench.py's CPU arm times."""
    valid = cls >= 0
    n_valid = valid.sum()
    in_sample = None
    if query_idx is not None:
        in_sample = torch.zeros_like(valid)
        in_sample[query_idx] = True
    total = ehat.new_zeros(())
    for c in torch.unique(cls[valid]).tolist():
        sel = cls == c
        q_sel = sel if in_sample is None else (sel & in_sample)
        if not bool(q_sel.any()):
            continue
        neg = valid & ~sel
        total = total + info_nce(ehat[q_sel], ehat[sel], ehat[neg], temperature, reduction="none").sum()
    return total / n_valid


def dense_infonce_from_features(x, cls, temperature=0.1, num_classes=None):
    """Full op as the product exposes it: x [N,D] raw embeddings -> normalise -> dense_infonce."""
    return dense_infonce(l2_normalize_pixels(x), cls, temperature=temperature, num_classes=num_classes)


# --------------------------------------------------------------------------
# a2  image-level contrast (IMC)
# --------------------------------------------------------------------------


def image_level_contrast(emb, label):
    """src/loss_multilabel.py:36-66.

    e = normalize(emb, eps=1e-6) (:41).  For image i only partners j>i count (:46).
    positive  <=> multi-hot labels identical (:51); negative <=> labels disjoint (:54).
    sums are seeded with 1e-6 (:44-45); image i contributes -log(pos/(pos+neg)) only
    if it has >=1 positive, >=1 negative and more negatives than positives (:57);
    total / B (:65).  Returns python 0 when no image qualifies (caller checks
    torch.is_tensor, train_mcl.py:194)."""
    b = emb.shape[0]
    e = emb / emb.norm(dim=-1, keepdim=True).clamp_min(1e-6)
    sim = torch.exp((e @ e.t()) / 0.1)
    lab = label.long()
    same = (label[:, None, :] == label[None, :, :]).all(-1)
    disjoint = (lab[:, None, :] & lab[None, :, :]).sum(-1) == 0
    upper = torch.triu(torch.ones(b, b, dtype=torch.bool), diagonal=1)
    pos_m = same & upper
    neg_m = disjoint & upper
    n_pos = pos_m.sum(1)
    n_neg = neg_m.sum(1)
    s_pos = 1e-6 + (sim * pos_m.to(sim.dtype)).sum(1)
    s_neg = 1e-6 + (sim * neg_m.to(sim.dtype)).sum(1)
    use = (n_pos > 0) & (n_neg > 0) & (n_neg > n_pos)
    if not bool(use.any()):
        return 0
    terms = -torch.log(s_pos / (s_pos + s_neg))
    return (terms * use.to(sim.dtype)).sum() / b


# --------------------------------------------------------------------------
# a3  PixPro
# --------------------------------------------------------------------------


def _cosine_dim0(a, b, eps=1e-8):
    """F.cosine_similarity(a, b, dim=0) as torch 2.11 computes it: each norm clamped
    to eps separately."""
    na = a.norm(dim=0).clamp_min(eps)
    nb = b.norm(dim=0).clamp_min(eps)
    return (a * b).sum(0) / (na * nb)


def pixpro(fm1s, fm2s, coord1s, coord2s):
    """src/loss_multilabel.py:93-105.  box = [top, left, h, w]; view 2 is detached (:103)."""
    coord1s = coord1s.long()
    coord2s = coord2s.long()
    acc = 0
    for b in range(fm1s.shape[0]):
        t1, l1, h1, w1 = [int(v) for v in coord1s[b]]
        t2, l2, h2, w2 = [int(v) for v in coord2s[b]]
        a = fm1s[b, :, t1:t1 + h1, l1:l1 + w1]
        c = fm2s[b, :, t2:t2 + h2, l2:l2 + w2].detach()
        acc = acc + _cosine_dim0(a, c).mean()
    return 1 - acc / fm1s.shape[0]


# --------------------------------------------------------------------------
# a4-a8  region EMD: crops feeder, cost matrix, weights, log-domain Sinkhorn
# --------------------------------------------------------------------------


def pair_wise_cos_cost(x_flat, y_flat):
    """src/loss_multilabel.py:232-242 with form='CV': x [N,C], y [M,C] -> 1 - x y^T  [N,M]."""
    return 1 - x_flat @ y_flat.t()


def region_weight(a_flat, b_flat):
    """src/loss_multilabel.py:250-257: every pixel of A dotted with the mean pixel of B.
    a_flat [Na,C], b_flat [Nb,C] -> [Na]."""
    return a_flat @ b_flat.mean(0)


def sinkhorn_logsumexp(cost, w1, w2, reg=0.1, maxiter=10):
    """src/loss_multilabel.py:201-230.

    M = (-C + u_i + v_j)/reg (:204);  per iteration BOTH updates use the same
    pre-update M (:215-217);  after the loop pi = exp(M) and the result is
    sum(pi * C.detach()) / numel (:219-221)."""
    u = torch.zeros_like(w1)
    v = torch.zeros_like(w2)
    log_mu = torch.log(w1 + 1e-6)
    log_nu = torch.log(w2 + 1e-6)
    for _ in range(maxiter):
        m = (-cost + u[:, None] + v[None, :]) / reg
        u = reg * (log_mu - torch.logsumexp(m, dim=1)) + u
        v = reg * (log_nu - torch.logsumexp(m, dim=0)) + v
    m = (-cost + u[:, None] + v[None, :]) / reg
    return (torch.exp(m) * cost.detach()).sum() / cost.numel()


def _flat_pixels(crop):
    """[1,C,h,w] -> [h*w, C]  (src/loss_multilabel.py:300,306)."""
    return crop.reshape(crop.shape[1], -1).t()


def emd_dynamic(crops1, crops2):
    """src/loss_multilabel.py:287-326 (EMD()(…, mode='dynamic')).

    Per image: score every (crop1, crop2) pair without grad (:308-313), keep the pair
    with the smallest score (:318 — dict keyed by 0-dim tensors, so equal scores do not
    collapse; the stable sort keeps the first), recompute cost + Sinkhorn WITH grad using
    the no-grad weights (:322-323).  Mean over images (:326)."""
    total = 0
    for img_crops1, img_crops2 in zip(crops1, crops2):
        best = None
        with torch.no_grad():
            for c1 in img_crops1:
                x = _flat_pixels(c1)
                for c2 in img_crops2:
                    y = _flat_pixels(c2)
                    w1 = region_weight(x, y)
                    w2 = region_weight(y, x)
                    score = float(sinkhorn_logsumexp(pair_wise_cos_cost(x, y), w1, w2))
                    if best is None or score < best[0]:
                        best = (score, c1, c2, w1, w2)
        _, c1, c2, w1, w2 = best
        cost = pair_wise_cos_cost(_flat_pixels(c1), _flat_pixels(c2))
        total = total + sinkhorn_logsumexp(cost, w1, w2)
    return total / len(crops1)


def static_target_shape(h, w, long_side=7):
    """src/loss_multilabel.py:265-269 (and :275-278): the output size handed to F.interpolate —
    (round(w*7/h), 7) when w < h, else (7, round(h*7/w)); python round (half to even)."""
    if w < h:
        return (int(round(w * long_side / h)), long_side)
    return (long_side, int(round(h * long_side / w)))


def emd_static(crops1, crops2):
    """src/loss_multilabel.py:259-285 (EMD()(…, mode='static'), the default mode).

    Pair i: both regions resized (bilinear, align_corners=True) to static_target_shape, cost
    1 - x y^T over all position pairs and the batch dim (:281), its mean (:282); region 2 is
    resized without grad (:273).  The list of means is sorted and the two smallest are summed
    (:283-284).  Returns (loss, dists [L])."""
    dists = []
    for a, b in zip(crops1, crops2):
        n, c = a.shape[:2]
        x = F.interpolate(a, size=static_target_shape(*a.shape[2:]), mode="bilinear", align_corners=True)
        with torch.no_grad():
            y = F.interpolate(b, size=static_target_shape(*b.shape[2:]), mode="bilinear", align_corners=True)
        x, y = x.reshape(n, c, -1), y.reshape(n, c, -1)
        cost = 1 - torch.einsum("bcn,bcm->bnm", x, y)
        dists.append(cost.mean())
    order = sorted(range(len(dists)), key=lambda i: float(dists[i].detach()))      # stable, like list.sort()
    return dists[order[0]] + dists[order[1]], torch.stack([d.detach() for d in dists])


def get_static_crops(x1, coord1, x2, coord2):
    """src/torchutils.py:177-214: per image the overlap boxes (y0, x0, h, w) of both views;
    boxes under 15 px or more elongated than 7:1 are skipped (:199); view 1 is cut into nine
    half-size windows at offsets (0, h//4, h//2) x (0, w//4, w//2) (:203-205), view 2 into nine
    third-size windows at (0, h//3, 2h//3) x (0, w//3, 2w//3) (:210-213), all with view 1's h, w.
    Returns flat lists of [1,C,·,·] views and the image index of every window."""
    crops1, crops2, batch_indices = [], [], []
    for b in range(x1.shape[0]):
        b1, b2 = [int(v) for v in coord1[b]], [int(v) for v in coord2[b]]
        ins1 = x1[b:b + 1, :, b1[0]:b1[0] + b1[2], b1[1]:b1[1] + b1[3]]
        ins2 = x2[b:b + 1, :, b2[0]:b2[0] + b2[2], b2[1]:b2[1] + b2[3]]
        h, w = ins1.shape[2:]
        if h < 15 or w < 15 or h / w > 7 or w / h > 7:
            continue
        for i in (0, h // 4, h // 2):
            for j in (0, w // 4, w // 2):
                crops1.append(ins1[:, :, i:i + h // 2, j:j + w // 2])
        for i in (0, h // 3, 2 * h // 3):
            for j in (0, w // 3, 2 * w // 3):
                crops2.append(ins2[:, :, i:i + h // 3, j:j + w // 3])
                batch_indices.append(b)
    return crops1, crops2, batch_indices


def dynamic_crop_plan(h, w, rng=np.random):
    """Integer half of src/torchutils.py:240-280 for ONE overlap box of size h x w:
    returns None when the box is rejected (:240), else
    (view-1 windows [(top,left,lh,lw)], resized (rh,rw), pool4?, view-2 windows [(top,left,hh,ww)]).
    Consumes the numpy RNG exactly like the reference (:246-254)."""
    if h < 15 or w < 15 or h / w > 5 or w / h > 5:
        return None
    lh = rng.randint(h // 3, h // 2 + 1)
    lw = rng.randint(w // 3, w // 2 + 1)
    while lh < 5 or lw < 5:
        lh = rng.randint(h // 3, h // 2 + 1)
        lw = rng.randint(w // 3, w // 2 + 1)
    sh = rng.randint(lh // 2, lh + 1)
    sw = rng.randint(lw // 2, lw + 1)
    rh = round(h / (h / sh))
    rw = round(w / (w / sw))
    wins1 = []
    if rh >= 7 and rw >= 7:
        for top in range(0, h, sh):
            for left in range(0, w, sw):
                if top + lh <= h and left + lw <= w:
                    wins1.append((top, left, lh, lw))
    wins2 = [(top, left, h // 2, w // 2)
             for top in range(0, h - 1, h // 2) for left in range(0, w - 1, w // 2)]
    return wins1, (rh, rw), (rh > 28 or rw > 28), wins2


def get_dynamic_crops(x1, coord1, x2, coord2, rng=np.random):
    """src/torchutils.py:217-291: overlap boxes -> random windows of view 1 resized
    bilinearly (align_corners=True, :270; 4x4 avg-pool if a side > 28, :271-272) and the
    2x2 (up to 3x3 for odd sizes, :276-277) grid of view 2 avg-pooled 4x4 (:279)."""
    crops1, crops2, batch_indices = [], [], []
    for b in range(x1.shape[0]):
        t1, l1, h, w = [int(v) for v in coord1[b]]
        t2, l2, h2, w2 = [int(v) for v in coord2[b]]
        box1 = x1[b:b + 1, :, t1:t1 + h, l1:l1 + w]
        box2 = x2[b:b + 1, :, t2:t2 + h2, l2:l2 + w2]
        h, w = box1.shape[2], box1.shape[3]
        plan = dynamic_crop_plan(h, w, rng)
        if plan is None:
            continue
        wins1, (rh, rw), pool, wins2 = plan
        img1 = []
        for top, left, lh, lw in wins1:
            patch = F.interpolate(box1[:, :, top:top + lh, left:left + lw], (rh, rw),
                                  mode="bilinear", align_corners=True)
            if pool:
                patch = F.avg_pool2d(patch, 4, 4)
            img1.append(patch)
        img2 = [F.avg_pool2d(box2[:, :, top:top + hh, left:left + ww], 4, 4)
                for top, left, hh, ww in wins2]
        if img1:
            batch_indices.append(b)
            crops1.append(img1)
            crops2.append(img2)
    return crops1, crops2, batch_indices


# --------------------------------------------------------------------------
# a9-a11  CAM head, image embedding/logits, projection + PCM
# --------------------------------------------------------------------------


def cam_head(p7, fc_weight):
    """src/MuSCLe.py:242-247 (== :263-268): relu(sum_ch p7 * W) — a 1x1 conv with
    W = fc.weight.data, i.e. no gradient reaches W."""
    return torch.relu(torch.einsum("bchw,kc->bkhw", p7, fc_weight.detach()))


def image_embedding_logits(p7, fc_weight):
    """src/MuSCLe.py:182-183,240-241: emb = GAP(p7); logits = fc(emb) (no bias; grad reaches W)."""
    emb = p7.mean(dim=(2, 3))
    return emb, emb @ fc_weight.t()


def pcm(cam, f, fuse_weight, fuse_bias):
    """src/MuSCLe.py:213-223.

    g = fuse(f) (1x1 conv, :216);  ghat = g / (||g||_2 + 1e-5) over channels (:218);
    S = relu(ghat^T ghat) [hw,hw] (:220);  S /= (column sums + 1e-5) (:221);
    out = cam @ S (:222).  cam is resized bilinearly to f's size first (:215; identity
    when the sizes already agree)."""
    n, _, h, w = f.shape
    if cam.shape[2:] != (h, w):
        cam = F.interpolate(cam, (h, w), mode="bilinear", align_corners=True)
    g = torch.einsum("bchw,kc->bkhw", f, fuse_weight.reshape(fuse_weight.shape[0], -1))
    g = g + fuse_bias[None, :, None, None]
    g = g.reshape(n, g.shape[1], h * w)
    ghat = g / (g.norm(dim=1, keepdim=True) + 1e-5)
    s = torch.relu(ghat.transpose(1, 2) @ ghat)
    s = s / (s.sum(dim=1, keepdim=True) + 1e-5)
    out = cam.reshape(n, cam.shape[1], h * w) @ s
    return out.reshape(n, cam.shape[1], h, w)


# --------------------------------------------------------------------------
# a18  the trainers' CAM normalisers
# --------------------------------------------------------------------------


def cam_softmaxnorm(cams):
    """train_mcl.py:30-36: softmax over the fg channels, bg = 1 - max fg."""
    fg = torch.softmax(cams[:, 1:], dim=1)
    bg = 1 - fg.max(dim=1, keepdim=True)[0]
    return torch.cat([bg, fg], dim=1)


def cam_maxnorm(cams):
    """train_mcl.py:21-28: relu, per-(image,class) min-max with -1e-6 / +1e-6, relu."""
    cams = torch.relu(cams)
    flat = cams.flatten(2)
    lo = flat.min(-1)[0][:, :, None, None]
    hi = flat.max(-1)[0][:, :, None, None]
    return torch.relu((cams - lo - 1e-6) / (hi - lo + 1e-6))


# --------------------------------------------------------------------------
# a12-a17  FieldLoss (boundary-aware pseudo-mask pixel-pixel contrast)
# --------------------------------------------------------------------------

_SOBEL5_X = [[2.0, 1.0, 1e-6, -1.0, -2.0],
             [3.0, 2.0, 1e-6, -2.0, -3.0],
             [4.0, 3.0, 0.0, -3.0, -4.0],
             [3.0, 2.0, 1e-6, -2.0, -3.0],
             [2.0, 1.0, 1e-6, -1.0, -2.0]]
_SOBEL5_Y = [[2.0, 3.0, 4.0, 3.0, 2.0],
             [1.0, 2.0, 3.0, 2.0, 1.0],
             [1e-6, 1e-6, 1e-6, 1e-6, 1e-6],
             [-1.0, -2.0, -3.0, -2.0, -1.0],
             [-2.0, -3.0, -4.0, -3.0, -2.0]]
_SOBEL3_X = [[1.0, 1e-6, -1.0], [2.0, 1e-6, -2.0], [1.0, 1e-6, -1.0]]
_SOBEL3_Y = [[1.0, 2.0, 1.0], [1e-6, 1e-6, 1e-6], [-1.0, -2.0, -1.0]]


def sobel_kernels(size, dtype=torch.float32):
    """src/edge.py:33-43 (values incl. the 1e-6 'zeros' and the exact 0.0 centre of Gx5)."""
    gx, gy = (_SOBEL3_X, _SOBEL3_Y) if size == 3 else (_SOBEL5_X, _SOBEL5_Y)
    # the reference builds the filters in float32, then convolves in the input dtype
    k = torch.tensor([gx, gy], dtype=torch.float32).to(dtype)
    return k[:, None]  # [2,1,s,s]


def class_edges(seg_map, label_with_bg, beta=100.0, sobel_size=5):
    """src/edge.py:423-440 (Mix_fg.forward, reduction=False): softmax(seg*beta) over
    classes (:425), drop bg (:426), Sobel x/y per fg class (:429-431, cross-correlation,
    zero padding), zero the classes absent from the image labels (:437).
    -> [n, C-1, 2, h, w]."""
    n, c, h, w = seg_map.shape
    prob = torch.softmax(seg_map * beta, dim=1)[:, 1:]
    k = sobel_kernels(sobel_size, seg_map.dtype)
    g = F.conv2d(prob.reshape(n * (c - 1), 1, h, w), k, padding=sobel_size // 2)
    g = g.reshape(n, c - 1, 2, h, w)
    return g * label_with_bg[:, 1:, None, None, None].to(g.dtype)


_ORIENT_DIV = 3.1416 / 8  # src/edge.py:63


def orient_quantize(edges):
    """src/edge.py:65-89 for a [n,C-1,2,h,w] input: magnitude sqrt(gx^2+gy^2+1e-8) (:70) and
    the 8-bin quantisation of atan2(gy,gx) with bin edges at odd multiples of 3.1416/8
    (:72-88).  Returned bins are 0..7 as floats, like the reference."""
    mag = torch.sqrt((edges ** 2).sum(dim=2) + 1e-8)
    ang = torch.atan2(edges[:, :, 1], edges[:, :, 0])
    d = _ORIENT_DIV
    bins = torch.full_like(ang, 3.0)              # |ang| >= 7d  (mask4, :75-76)
    lo_hi_bin = [(1, 3, 0), (3, 5, 1), (5, 7, 2), (-7, -5, 4), (-5, -3, 5), (-3, -1, 6), (-1, 1, 7)]
    for lo, hi, b in lo_hi_bin:
        bins = torch.where((ang >= lo * d) & (ang < hi * d), torch.full_like(ang, float(b)), bins)
    return mag, bins


def boundary_positions(mag, label_with_bg):
    """src/edge.py:370-374: per (image,class) keep pixels with mag >= 0.8*max where
    max > 1 and the class is present."""
    n, c = mag.shape[:2]
    mx = mag.reshape(n, c, -1).max(-1)[0][:, :, None, None]
    pos = (mag >= 0.8 * mx) & (mx > 1)
    return pos & label_with_bg[:, 1:, None, None].bool()


def in_out_indices(orient_plus1, pos, step):
    """src/edge.py:195-227 (single step list, as called at :288).

    orient_plus1 holds 1..8.  Offsets of the reference's float formula:
      4, 8  -> outside = idx - step*w, inside = idx + w
      2, 6  -> outside = idx - 1,      inside = idx + 1
      odd   -> outside = inside = idx
    then indices with idx % (w-1) in {0,1}, idx <= 0 or idx >= w*h-1 are dropped (:220-221)."""
    h, w = orient_plus1.shape
    idx = torch.arange(h * w, dtype=torch.int64).reshape(h, w)
    o = orient_plus1.to(torch.int64)
    vert = (o % 4 == 0)
    horiz = (o == 2) | (o == 6)
    out_all = idx - step * w * vert.to(torch.int64) - horiz.to(torch.int64)
    in_all = idx + w * vert.to(torch.int64) + horiz.to(torch.int64)
    out_i = out_all[pos]
    in_i = in_all[pos]

    def keep(t):
        return (t % (w - 1) != 0) & (t % (w - 1) != 1) & (t > 0) & (t < w * h - 1)

    return out_i[keep(out_i)], in_i[keep(in_i)]


def beacon_terms(fp, fn, tp, tn, sim, dim):
    """src/edge.py:229-258: -mean(FP) + mean(FN) + mean(TP) - mean(TN) of sim.mean(dim),
    each term only when its set is non-empty."""
    m = sim.mean(dim)
    loss = 0
    for sel, sign in ((fp, -1.0), (fn, 1.0), (tp, 1.0), (tn, -1.0)):
        if bool(sel.any()):
            loss = loss + sign * m[sel].mean()
    return loss


def fieldloss_bifilter(orient, dense_ft, mask, pos_idx, label_fg, step, k, rng=_pyrandom):
    """src/edge.py:261-361.  Channel softmax of dense_ft and mask (:271-272); per
    (image, present class): in/out index lists (:288), gather columns (:290-294);
    if both lists are longer than k draw k of each with random.sample — outside first,
    then inside (:297-304); otherwise the pair is skipped (:306-316);
    sim_out = outs^T @ ins.detach(), sim_mask likewise on the pseudo-mask (:318-319);
    'sign' = row/column mean above the global mean (:321-322,332-333) -> FP/FN/TP/TN
    (:323-326,335-338) -> beacon_terms on rows then columns (:328-341);
    result loss/n (:359-361; python 0 if nothing contributed)."""
    n, ch, h, w = dense_ft.shape
    dsm = torch.softmax(dense_ft, dim=1).reshape(n, ch, -1)
    msm = torch.softmax(mask, dim=1).reshape(n, mask.shape[1], -1)
    loss = 0
    for b in range(n):
        for c in range(orient.shape[1]):
            if label_fg[b, c] == 0:
                continue
            with torch.no_grad():
                out_i, in_i = in_out_indices(orient[b, c] + 1, pos_idx[b, c], step)
            if not (in_i.numel() > k and out_i.numel() > k):
                continue
            sel_out = rng.sample(range(out_i.numel()), k)
            sel_in = rng.sample(range(in_i.numel()), k)
            out_i = out_i[sel_out]
            in_i = in_i[sel_in]
            outs, ins = dsm[b][:, out_i], dsm[b][:, in_i]
            outs_m, ins_m = msm[b][:, out_i].detach(), msm[b][:, in_i].detach()
            sim_out = outs.t() @ ins.detach()
            sim_mask = outs_m.t() @ ins_m
            for dim in (1, 0):
                s_mask = sim_mask.mean(dim) > sim_mask.mean()
                s_dense = (sim_out.mean(dim) > sim_out.mean()).detach()
                loss = loss + beacon_terms(s_mask & ~s_dense, ~s_mask & s_dense,
                                           ~s_mask & ~s_dense, s_mask & s_dense, sim_out, dim)
    if torch.is_tensor(loss):
        return loss / n
    return loss


def fieldloss_forward(seg_map, dense_ft, mask, label_with_bg, step=7, k=100, beta=1e3,
                      sobel_size=5, rng=_pyrandom):
    """src/edge.py:364-384: (loss | False, mag_fg.sum(1)).  The bifilter only runs when at
    least 10 boundary pixels were found (:378)."""
    with torch.no_grad():
        edges = class_edges(seg_map, label_with_bg, beta=beta, sobel_size=sobel_size)
        mag, orient = orient_quantize(edges)
        pos = boundary_positions(mag, label_with_bg)
        pos_count = int(pos.sum())
        mag_sum = mag.sum(1)
    loss = False
    if pos_count >= 10:
        loss = fieldloss_bifilter(orient, dense_ft, mask, pos, label_with_bg[:, 1:], step, k, rng)
    return loss, mag_sum


# --------------------------------------------------------------------------
# f1  ER loss and the bilinear up-sampling in front of it
# --------------------------------------------------------------------------


def upsample_bilinear(x, size):
    """src/MuSCLe.py:256-257: F.interpolate(x, size=(H,W), mode='bilinear', align_corners=True).
    Source coordinate of output pixel o along an axis of n_in -> n_out samples: o*(n_in-1)/(n_out-1)
    (0 when n_out == 1); linear weights between floor and floor+1 (clamped)."""
    return F.interpolate(x, size=tuple(size), mode="bilinear", align_corners=True)


def er_loss(raw_cams, raw_sgcs, label):
    """train_mcl.py:175-188: cams = cam_softmaxnorm(raw_cams).detach(), sgcs = cam_softmaxnorm(raw_sgcs) (:175-176);
    both masked by the image labels incl. background (:185-186); per image the k = int(0.2 * valid_channel * h * w)
    largest |cams - sgcs| over (class, pixel) with valid_channel = int(label.sum()) over the WHOLE batch (:178),
    and the mean of all n*k values (:188).  Gradient reaches raw_sgcs only."""
    n, c, h, w = raw_cams.shape
    label_with_bg = torch.cat([torch.ones(n, 1, dtype=label.dtype), label], dim=1)
    cams = cam_softmaxnorm(raw_cams).detach() * label_with_bg[:, :, None, None]
    sgcs = cam_softmaxnorm(raw_sgcs) * label_with_bg[:, :, None, None]
    valid_channel = int(label.sum())
    k = int(0.2 * valid_channel * h * w)
    diff = torch.abs(cams - sgcs).flatten(1)
    return torch.topk(diff, k=k, dim=-1)[0].mean()


# --------------------------------------------------------------------------
# f4  classification losses of the MCL stage
# --------------------------------------------------------------------------


def focal_loss(prob, target, gamma=2, alpha=0.5):
    """src/loss_multilabel.py:68-91 (FocalLoss.forward): pt = p if y else 1-p; -alpha (1-pt)^gamma log(pt + 1e-9),
    summed over classes, mean over images."""
    pt = target * prob + (1 - target) * (1 - prob)
    focal = -alpha * (1.0 - pt) ** gamma * torch.log(pt + 1e-9)
    return focal.sum(1).mean()


def lsep_loss(pred, labels):
    """src/loss_multilabel.py:24-33 (Log_Sum_Exp_Pairwise_Loss): pos = pred where label == 1 else 0, neg = pred where
    label == 0 else 0 (so the zeros take part in the pairs); log(1 + mean_{i,j} exp(neg_j - pos_i)) per image -> [N]."""
    pos = torch.where(labels == 0, torch.zeros_like(pred), pred)
    neg = torch.where(labels == 1, torch.zeros_like(pred), pred)
    exp_sub = torch.exp(neg[:, None, :] - pos[:, :, None])
    return torch.log(1 + exp_sub.sum(dim=(1, 2)) / (exp_sub.shape[1] * exp_sub.shape[2]))


def soft_margin_loss(logits, target):
    """nn.MultiLabelSoftMarginLoss (train_mcl.py:144,181): mean over images of the class-mean of
    -(y log sigmoid(x) + (1-y) log sigmoid(-x))."""
    per = -(target * F.logsigmoid(logits) + (1 - target) * F.logsigmoid(-logits))
    return per.mean(1).mean()


def mcl_cls_loss(logits, label):
    """train_mcl.py:180-183: focal(sigmoid(l[:,1:])) + soft margin(l[:,1:]) + LSEP(sigmoid(l[:,1:])).mean().
    -> (loss_cls, loss_focal, loss_softmargin, loss_pair)."""
    fg = logits[:, 1:]
    prob = torch.sigmoid(fg)
    lf = focal_loss(prob, label)
    ls = soft_margin_loss(fg, label)
    lp = lsep_loss(prob, label).mean()
    return lp + ls + lf, lf, ls, lp


# --------------------------------------------------------------------------
# f2  decoder tail feeding FieldLoss
# --------------------------------------------------------------------------


def decoder_tail(p3_dec, fuse_w, fuse_b, size):
    """src/MuSCLe.py:285-286: dense_ft = bilinear up-sampling of p3_dec to (H, W) (align_corners=True);
    seg_map = fuse_dec(dense_ft), a 1x1 convolution with bias.  -> (seg_map, dense_ft)."""
    dense_ft = upsample_bilinear(p3_dec, size)
    seg_map = torch.einsum("kc,nchw->nkhw", fuse_w.reshape(fuse_w.shape[0], -1), dense_ft) + fuse_b[None, :, None, None]
    return seg_map, dense_ft


# --------------------------------------------------------------------------
# f3  IRN random walk
# --------------------------------------------------------------------------


def to_transition_matrix(affinity_dense, beta, times):
    """src/indexing.py:116-123: element-wise power beta, columns normalised to sum 1, then squared `times` times
    (T <- T @ T), i.e. the 2^times-step transition matrix."""
    scaled = torch.pow(affinity_dense, beta)
    trans = scaled / scaled.sum(dim=0, keepdim=True)
    for _ in range(times):
        trans = trans @ trans
    return trans


def random_walk(x, edge, trans_mat):
    """src/indexing.py:144-148: rw = (x * (1 - edge)).view(-1, h*w) @ trans_mat -> [C, 1, h, w]."""
    h, w = x.shape[-2:]
    xx = x.reshape(-1, h, w) * (1 - edge)
    rw = xx.reshape(-1, h * w) @ trans_mat
    return rw.reshape(rw.shape[0], 1, h, w)


# --------------------------------------------------------------------------
# synthetic inputs of the tests: the generator lives beside bench.py (workloads.py); these keep the
# tests' historical inputs (background channel biased by +0.15, i.e. ~75 % background pixels)
# --------------------------------------------------------------------------
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import workloads as _workloads  # noqa: E402

VOC_CLASS_FREQ = _workloads.VOC_CLASS_FREQ
synth_labels = _workloads.synth_labels


def synth_pixel_classes(label, h, w, gen, smooth=9, bg_bias=0.15):
    return _workloads.synth_pixel_classes(label, h, w, gen, smooth=smooth, bg_bias=bg_bias)


def synth_embeddings(batch, dim, h, w, seed, bg_bias=0.15):
    """(x [B,D,h,w] fp32, cls [B,h,w] int64, label [B,20]); see workloads.synth_embeddings."""
    return _workloads.synth_embeddings(batch, dim, h, w, seed, bg_bias=bg_bias)
