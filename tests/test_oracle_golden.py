"""The oracle (oracle/muscle_oracle.py) against fixtures minted by the unmodified reference
(oracle/make_golden.py).  CPU only."""
import random

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O


def close(a, b, rtol=1e-5, atol=1e-6):
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("tag", ["f32", "f64"])
def test_info_nce(golden, tag):
    g = golden("info_nce")[tag]
    q, p, k = (g[n].clone().requires_grad_() for n in "qpk")
    loss = O.info_nce(q, p, k)
    gq, gp, gk = torch.autograd.grad(loss, (q, p, k))
    tol = dict(rtol=1e-5, atol=1e-6) if tag == "f32" else dict(rtol=1e-11, atol=1e-12)
    close(loss, g["loss"], **tol)
    close(gq, g["gq"], **tol)
    close(gp, g["gp"], **tol)
    close(gk, g["gk"], **tol)
    close(O.info_nce(q, p, k, temperature=0.07, reduction="none"), g["none_t007"], **tol)


@pytest.mark.parametrize("tag", ["f32", "f64"])
def test_info_nce_rows_of_any_norm(golden, tag):
    """the primitive never normalises: rows of any length, reduction='none' with an upstream gradient per row"""
    g = golden("info_nce_anynorm")[tag]
    q, p, k = (g[n].clone().requires_grad_() for n in "qpk")
    tol = dict(rtol=2e-5, atol=1e-6) if tag == "f32" else dict(rtol=1e-11, atol=1e-12)
    loss = O.info_nce(q, p, k, temperature=0.3)
    for got, want in zip(torch.autograd.grad(loss, (q, p, k)), (g["gq"], g["gp"], g["gk"])):
        close(got, want, **tol)
    close(loss, g["loss"], **tol)
    rows = O.info_nce(q, p, k, temperature=0.3, reduction="none")
    close(rows, g["rows"], **tol)
    for got, want in zip(torch.autograd.grad((rows * g["up"]).sum(), (q, p, k)), (g["nq"], g["np"], g["nk"])):
        close(got, want, **tol)


@pytest.mark.parametrize("tag", ["f32", "f64"])
def test_dense_infonce_composition(golden, tag):
    g = golden("dense_infonce")[tag]
    x = g["x"].clone().requires_grad_()
    loss = O.dense_infonce_from_features(x, g["cls"], temperature=0.1, num_classes=5)
    gx, = torch.autograd.grad(loss, x)
    tol = dict(rtol=2e-5, atol=1e-6) if tag == "f32" else dict(rtol=1e-10, atol=1e-12)
    close(loss, g["loss"], **tol)
    close(gx, g["gx"], **tol)


def test_dense_infonce_chunking_and_anchor(golden):
    g = golden("dense_infonce")["f64"]
    e = O.l2_normalize_pixels(g["x"])
    a = O.dense_infonce(e, g["cls"], row_chunk=7)
    b = O.dense_infonce(e, g["cls"], row_chunk=4096)
    close(a, b, rtol=1e-12, atol=0)
    # hard-mask anchor: 0/1 weights as "CAM weights" give the same prototypes as the mean
    valid = g["cls"] >= 0
    w = torch.zeros(e.shape[0], 5, dtype=e.dtype)
    w[valid, g["cls"][valid]] = 1
    p_soft, _ = O.class_prototypes(e, g["cls"], 5, weights=w)
    p_hard, _ = O.class_prototypes(e, g["cls"], 5)
    close(p_soft, p_hard, rtol=1e-12, atol=0)


def test_image_level_contrast(golden):
    g = golden("imc")
    emb = g["emb"].clone().requires_grad_()
    loss = O.image_level_contrast(emb, g["label"])
    ge, = torch.autograd.grad(loss, emb)
    close(loss, g["loss"], rtol=1e-5, atol=1e-6)
    close(ge, g["ge"], rtol=1e-4, atol=1e-6)
    label0 = torch.zeros(3, 20)
    label0[:, 1] = 1
    z = O.image_level_contrast(torch.randn(3, 48), label0)
    assert not torch.is_tensor(z) and z == 0 and not g["zero_is_tensor"]
    emb2 = g["emb2"].clone().requires_grad_()
    loss2 = O.image_level_contrast(emb2, g["label2"])
    if g["loss2"] is None:
        assert not torch.is_tensor(loss2)
    else:
        close(loss2, g["loss2"], rtol=1e-5, atol=1e-6)
        close(torch.autograd.grad(loss2, emb2)[0], g["ge2"], rtol=1e-4, atol=1e-6)


def test_pixpro(golden):
    g = golden("pixpro")
    fm1 = g["fm1"].clone().requires_grad_()
    loss = O.pixpro(fm1, g["fm2"], g["c1"], g["c2"])
    close(loss, g["loss"], rtol=1e-5, atol=1e-6)
    close(torch.autograd.grad(loss, fm1)[0], g["g1"], rtol=1e-4, atol=1e-7)


def test_emd_dynamic(golden):
    g = golden("emd")
    raw1 = g["raw1"].clone().requires_grad_()
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(g["raw2"] * 4, 1), dim=1)
    rng = np.random.RandomState(g["np_seed"])
    crops1, crops2, bidx = O.get_dynamic_crops(vw1, g["c1"], vw2.detach(), g["c2"], rng=rng)
    assert bidx == g["batch_indices"]
    assert [[tuple(t.shape) for t in c] for c in crops1] == g["crop_shapes1"]
    assert [[tuple(t.shape) for t in c] for c in crops2] == g["crop_shapes2"]
    loss = O.emd_dynamic(crops1, crops2)
    close(loss, g["loss"], rtol=1e-4, atol=1e-7)
    close(torch.autograd.grad(loss, raw1)[0], g["g1"], rtol=1e-3, atol=1e-8)
    sk = O.sinkhorn_logsumexp(g["sk_cost"], g["sk_w1"].squeeze(0), g["sk_w2"].squeeze(0))
    close(sk, g["sk_value"], rtol=1e-5, atol=1e-8)


def test_emd_static(golden):
    g = golden("emd_static")
    raw1 = g["raw1"].clone().requires_grad_()
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(g["raw2"] * 4, 1), dim=1)
    crops1, crops2, bidx = O.get_static_crops(vw1, g["c1"], vw2.detach(), g["c2"])
    assert bidx == g["batch_indices"] and len(crops1) == len(crops2) == 18
    assert [tuple(t.shape) for t in crops1] == g["crop_shapes1"]
    assert [tuple(t.shape) for t in crops2] == g["crop_shapes2"]
    loss, dists = O.emd_static(crops1, crops2)
    close(dists, g["dists"], rtol=1e-5, atol=1e-7)
    close(loss, g["loss"], rtol=1e-5, atol=1e-7)
    close(torch.autograd.grad(loss, raw1)[0], g["g1"], rtol=1e-4, atol=1e-9)


def test_heads_in_situ(golden):
    g = golden("heads")
    p1, p3, p5, p7 = g["p1"], g["p3"], g["p5"], g["p7"]
    emb, logits = O.image_embedding_logits(p7, g["fc_w"])
    close(emb, g["emb"], rtol=1e-5, atol=1e-6)
    close(logits, g["logits"], rtol=1e-4, atol=1e-5)
    cams = O.cam_head(p7, g["fc_w"])
    hw = p7.shape[2:]
    fs = torch.cat([torch.relu(F.interpolate(p1, size=hw, mode="bilinear", align_corners=True)),
                    torch.relu(F.interpolate(p3, size=hw, mode="bilinear", align_corners=True)),
                    torch.relu(p5)], dim=1)
    sgc = O.pcm(cams, fs, g["fuse_w"], g["fuse_b"])
    up = lambda t: F.interpolate(t, size=g["img_hw"], mode="bilinear", align_corners=True)
    close(up(cams), g["cams_up"], rtol=1e-4, atol=1e-5)
    close(up(sgc), g["sgc_up"], rtol=1e-4, atol=1e-5)


def test_pcm_gradients(golden):
    g = golden("heads")
    cam = g["pcm_cam"].clone().requires_grad_()
    w = g["fuse_w"].clone().requires_grad_()
    b = g["fuse_b"].clone().requires_grad_()
    out = O.pcm(cam, g["pcm_f"], w, b)
    close(out, g["pcm_out"], rtol=1e-4, atol=1e-6)
    (out * g["pcm_go"]).sum().backward()
    close(cam.grad, g["pcm_gcam"], rtol=1e-3, atol=1e-6)
    close(w.grad, g["pcm_gw"], rtol=2e-3, atol=2e-5)
    close(b.grad, g["pcm_gb"], rtol=2e-3, atol=2e-5)


def test_fieldloss_sobel3(golden):
    """the constructor's other filter size (src/edge.py:33-35)"""
    g = golden("fieldloss_sobel3")
    edges = O.class_edges(g["seg"], g["label_bg"], beta=g["beta"], sobel_size=3)
    _, orient = O.orient_quantize(edges)
    assert (orient.to(torch.uint8) == g["orient"]).float().mean() > 0.9999      # (atan2 of two conv2d results: same ops)
    dense = g["dense"].clone().requires_grad_()
    rng = random.Random(g["py_seed"])
    loss, mag_sum = O.fieldloss_forward(g["seg"], dense, g["mask"], g["label_bg"], step=g["step"], k=g["k"],
                                        beta=g["beta"], sobel_size=3, rng=rng)
    close(mag_sum, g["mag"], rtol=1e-4, atol=1e-4)
    close(loss, g["loss"], rtol=1e-4, atol=1e-7)
    close(torch.autograd.grad(loss, dense)[0], g["gd"], rtol=1e-3, atol=1e-9)


def test_fieldloss(golden):
    g = golden("fieldloss")
    edges = O.class_edges(g["seg"], g["label_bg"], beta=g["beta"], sobel_size=5)
    close(edges, g["edges"], rtol=1e-4, atol=1e-5)
    mag, orient = O.orient_quantize(g["edges"])
    assert torch.equal(orient, g["orient"])
    pos = O.boundary_positions(mag, g["label_bg"])
    assert torch.equal(pos, g["pos"])
    out_i, in_i = O.in_out_indices(g["orient"][0, 14] + 1, g["pos"][0, 14], g["step"])
    assert torch.equal(out_i, g["out_idx_0_14"]) and torch.equal(in_i, g["in_idx_0_14"])
    dense = g["dense"].clone().requires_grad_()
    rng = random.Random(g["py_seed"])
    loss, mag_sum = O.fieldloss_forward(g["seg"], dense, g["mask"], g["label_bg"], step=g["step"], k=g["k"],
                                        beta=g["beta"], rng=rng)
    close(mag_sum, g["mag"], rtol=1e-4, atol=1e-4)
    close(loss, g["loss"], rtol=1e-4, atol=1e-7)
    close(torch.autograd.grad(loss, dense)[0], g["gd"], rtol=1e-3, atol=1e-9)
    flat = torch.zeros(1, 21, 24, 24)
    only_bg = torch.cat([torch.ones(1, 1), torch.zeros(1, 20)], 1)
    loss0, _ = O.fieldloss_forward(flat, torch.randn(1, 24, 24, 24), flat, only_bg, step=3, k=12, beta=g["beta"])
    assert loss0 is False and g["gate_false"]
    loss1, _ = O.fieldloss_forward(g["seg"], g["dense"], g["mask"], g["label_bg"], step=g["step"], k=100000,
                                   beta=g["beta"])
    assert (not torch.is_tensor(loss1)) and loss1 == 0 and g["big_k_is_zero"]


@pytest.mark.parametrize("d", [64, 128])
def test_dense_infonce_composition_at_tensor_core_dims(golden, d):
    """fixtures of the dims the CUDA passes take: reference info_nce per class, fp32 and fp64"""
    g = golden("dense_infonce")["d%d" % d]
    for dt, tag, tol in ((torch.float32, "f32", dict(rtol=3e-5, atol=1e-6)), (torch.float64, "f64", dict(rtol=1e-10, atol=1e-12))):
        x = g["x"].to(dt).requires_grad_()
        loss = O.dense_infonce_from_features(x, g["cls"], temperature=0.1, num_classes=g["num_classes"])
        gx, = torch.autograd.grad(loss, x)
        close(loss, g["loss_" + tag], **tol)
        close(gx, g["gx_" + tag], **tol)
        e = O.l2_normalize_pixels(x.detach())
        close(O.dense_infonce_per_class(e, g["cls"], 0.1), g["loss_" + tag], **tol)


def test_cam_normalisers_pinned_to_the_reference(golden):
    """train_mcl.py:21-36 executed from the reference script (oracle/make_golden.py:gen_cam_norms)"""
    g = golden("cam_norms")
    for name, fn in (("maxnorm", O.cam_maxnorm), ("softmaxnorm", O.cam_softmaxnorm)):
        x = g["x"].clone().requires_grad_()
        y = fn(x)
        close(y, g[name]["y"], rtol=1e-6, atol=1e-7)
        gx, = torch.autograd.grad((y * g[name]["go"]).sum(), x)
        close(gx, g[name]["gx"], rtol=1e-5, atol=1e-6)


def test_er_loss_and_upsampling(golden):
    g = golden("er_loss")
    sg = g["sgc_small"].clone().requires_grad_()
    up_c = O.upsample_bilinear(g["cams_small"], g["size"])
    up_s = O.upsample_bilinear(sg, g["size"])
    close(up_c, g["up_cams"], rtol=1e-6, atol=1e-7)
    close(up_s, g["up_sgc"], rtol=1e-6, atol=1e-7)
    loss = O.er_loss(up_c, up_s, g["label"])
    close(loss, g["loss_er"], rtol=1e-6, atol=1e-8)
    gs, = torch.autograd.grad(loss, sg)
    close(gs, g["g_sgc_small"], rtol=1e-5, atol=1e-9)
    assert g["k"] == int(0.2 * g["valid_channel"] * g["size"][0] * g["size"][1])


def test_classification_losses(golden):
    g = golden("cls_losses")
    logits = g["logits"].clone().requires_grad_()
    loss, lf, ls, lp = O.mcl_cls_loss(logits, g["label"])
    close(lf, g["loss_focal"])
    close(ls, g["loss_softmargin"])
    close(lp, g["loss_pair"])
    close(loss, g["loss_cls"])
    close(torch.autograd.grad(loss, logits)[0], g["g_logits"], rtol=1e-5, atol=1e-7)
    close(O.lsep_loss(torch.sigmoid(g["logits"][:, 1:]), g["label"]), g["pair_rows"])


def test_decoder_tail_into_fieldloss(golden):
    g = golden("decoder_tail")
    p3 = g["p3_dec"].clone().requires_grad_()
    fw = g["fuse_w"].clone().requires_grad_()
    fb = g["fuse_b"].clone().requires_grad_()
    seg, dense = O.decoder_tail(p3, fw, fb, g["size"])
    close(seg, g["seg_map"], rtol=1e-4, atol=1e-4)      # |seg| reaches 120: fp32 summation order of the 1x1 conv
    close(dense, g["dense_ft"], rtol=1e-5, atol=1e-6)
    rng = random.Random(g["py_seed"])
    loss, mag = O.fieldloss_forward(seg, dense, g["mask"], g["label_bg"], step=g["step"], k=g["k"], beta=g["beta"], rng=rng)
    close(loss, g["loss"], rtol=1e-4, atol=1e-7)
    close(mag, g["mag"], rtol=3e-3, atol=3e-3)     # softmax(seg * 100) amplifies the 1e-5 of seg a hundredfold
    gp, = torch.autograd.grad(loss, p3, allow_unused=True)
    close(gp, g["g_p3"], rtol=2e-3, atol=1e-9)
    assert g["g_w"] is None and g["g_b"] is None      # seg_map enters FieldLoss under no_grad only (src/edge.py:366-374)


def test_irn_transition_matrix_and_random_walk(golden):
    g = golden("irn")
    for times in (1, 4, 8):
        close(O.to_transition_matrix(g["aff"], 10, times), g["trans_%d" % times], rtol=1e-4, atol=1e-9)
    close(O.to_transition_matrix(g["aff"], 2, 3), g["trans_b2_3"], rtol=1e-4, atol=1e-9)


def test_cam_normalisers():
    x = torch.randn(2, 21, 6, 7)
    sm = O.cam_softmaxnorm(x)
    assert sm.shape == x.shape
    close(sm[:, 1:].sum(1), torch.ones(2, 6, 7), rtol=1e-5, atol=1e-6)
    close(sm[:, 0], 1 - sm[:, 1:].max(1)[0])
    mn = O.cam_maxnorm(x)
    assert float(mn.min()) >= 0 and float(mn.max()) <= 1
