"""CAM-weighted prototype pooling and its gradient scatter (csrc/pool.cu; SURVEY.md §8b `pool_prototypes`, §8c soft
variant) against the oracle, against the reference's own instances of it (the positive-key mean of info_nce,
src/loss_multilabel.py:12, and `cam @ aff` of src/MuSCLe.py:222 via the reference-minted PCM fixture), and inside the
dense InfoNCE as soft positives."""
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


@pytest.mark.parametrize("n,d,c,dtype", [(4096, 128, 21, torch.float32), (5000, 128, 21, torch.bfloat16),
                                         (777, 64, 5, torch.float32), (3000, 256, 24, torch.float32), (130, 32, 1, torch.float32)])
def test_hard_mask_anchor(n, d, c, dtype):
    """0/1 weights: the prototype is the reference's positive_keys.mean (src/loss_multilabel.py:12) per class"""
    from muscle_b200 import pool_prototypes
    gen = torch.Generator().manual_seed(n + d)
    e = F.normalize(torch.randn(n, d, generator=gen), dim=1)
    cls = torch.randint(-1, c, (n,), generator=gen)
    w = torch.zeros(n, c)
    w[cls >= 0, cls[cls >= 0]] = 1
    eg = e.to(dtype).cuda()
    proto, sums, mass = pool_prototypes(eg, w.cuda(), return_all=True)
    e_ref = eg.float().cpu()
    for ci in range(c):
        sel = cls == ci
        assert float(mass[ci]) == float(sel.sum())
        if bool(sel.any()):
            torch.testing.assert_close(proto[ci].cpu(), e_ref[sel].mean(0), rtol=1e-4, atol=1e-6)


def test_cam_at_aff_is_an_instance(golden):
    """out = cam @ aff (src/MuSCLe.py:222): the un-normalised sums with e = rows of the normalised affinity, w = cam^T;
    the affinity comes from the oracle's restatement of :216-221, the expected output from the reference fixture"""
    from muscle_b200 import pool_prototypes
    g = golden("heads")
    cam, f = g["pcm_cam"], g["pcm_f"]
    n, _, h, w_ = f.shape
    gmap = torch.einsum("bchw,kc->bkhw", f, g["fuse_w"].reshape(g["fuse_w"].shape[0], -1)) + g["fuse_b"][None, :, None, None]
    gmap = gmap.reshape(n, -1, h * w_)
    ghat = gmap / (gmap.norm(dim=1, keepdim=True) + 1e-5)
    s = torch.relu(ghat.transpose(1, 2) @ ghat)
    s = s / (s.sum(dim=1, keepdim=True) + 1e-5)
    for b in range(n):
        _, sums, _ = pool_prototypes(s[b].contiguous().cuda(), cam[b].reshape(21, -1).t().contiguous().cuda(), return_all=True)
        ref = g["pcm_out"][b].reshape(21, -1)
        torch.testing.assert_close(sums.cpu(), ref, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("n,d,c,dtype", [(3000, 128, 21, torch.float32), (2048, 128, 21, torch.bfloat16), (1000, 64, 7, torch.float32),
                                         (65536, 128, 21, torch.bfloat16), (5000, 64, 7, torch.bfloat16),
                                         (300, 128, 24, torch.bfloat16), (33000, 256, 21, torch.bfloat16)])
def test_soft_weights_forward_and_scatter(n, d, c, dtype):
    from muscle_b200 import pool_prototypes
    gen = torch.Generator().manual_seed(n + c)
    e = F.normalize(torch.randn(n, d, generator=gen), dim=1).to(dtype)
    w = torch.relu(torch.randn(n, c, generator=gen))
    go = torch.randn(c, d, generator=gen)
    er = e.float().clone().requires_grad_()
    wr = w.clone().requires_grad_()
    pr, _ = O.class_prototypes(er, None, c, weights=wr)
    (pr * go).sum().backward()
    eg = e.cuda().requires_grad_()
    wg = w.cuda().requires_grad_()
    p = pool_prototypes(eg, wg)
    (p * go.cuda()).sum().backward()
    torch.testing.assert_close(p.detach().cpu(), pr.detach(), rtol=1e-4, atol=1e-6)
    assert _cos(wg.grad.cpu(), wr.grad) >= 0.99999
    torch.testing.assert_close(wg.grad.cpu(), wr.grad, rtol=1e-3, atol=1e-5 * float(wr.grad.abs().max()))
    assert _cos(eg.grad.float().cpu(), er.grad) >= (0.99999 if dtype == torch.float32 else 0.9999)
    if dtype == torch.float32:
        torch.testing.assert_close(eg.grad.cpu(), er.grad, rtol=1e-3, atol=1e-5 * float(er.grad.abs().max()))


def test_dense_infonce_with_cam_weighted_positives():
    """SURVEY §8c soft variant end to end: prototypes pooled with CAM-like weights are every pixel's positive; loss and
    the gradients w.r.t. the embeddings AND the weights against the oracle composition under autograd."""
    import muscle_b200 as M
    x, cls, _ = O.synth_embeddings(2, 128, 24, 24, seed=9)
    b, d, h, w_ = x.shape
    gen = torch.Generator().manual_seed(10)
    onehot = F.one_hot(cls.reshape(-1), 21).float()
    cam = torch.relu(0.7 * onehot + 0.3 * torch.rand(b * h * w_, 21, generator=gen) - 0.1)    # CAM scores around the mask
    flat = x.permute(0, 2, 3, 1).reshape(-1, d)
    # oracle
    xr = flat.clone().requires_grad_()
    wr = cam.clone().requires_grad_()
    er = O.l2_normalize_pixels(xr)
    proto_r, _ = O.class_prototypes(er, None, 21, weights=wr)
    ref = O.dense_infonce(er, cls.reshape(-1), temperature=0.1, num_classes=21, proto=proto_r)
    ref.backward()
    # CUDA path
    xg = flat.contiguous().cuda().requires_grad_()
    wg = cam.cuda().requires_grad_()
    eg = F.normalize(xg, dim=1)
    proto = M.pool_prototypes(eg, wg)
    loss = M.dense_infonce(xg, cls.reshape(-1).cuda(), 0.1, 21, proto=proto)
    loss.backward()
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item()), (loss.item(), ref.item())
    assert _cos(xg.grad.cpu(), xr.grad) >= 0.999
    assert _cos(wg.grad.cpu(), wr.grad) >= 0.999
    # hard anchor: one-hot weights reproduce the built-in class means (value), and the gradient of the modular path
    # (pool backward) equals the fused one
    xa = flat.contiguous().cuda().requires_grad_()
    la = M.dense_infonce(xa, cls.reshape(-1).cuda(), 0.1, 21)
    la.backward()
    xb = flat.contiguous().cuda().requires_grad_()
    pb = M.pool_prototypes(F.normalize(xb, dim=1), onehot.cuda())
    lb = M.dense_infonce(xb, cls.reshape(-1).cuda(), 0.1, 21, proto=pb)
    lb.backward()
    assert abs(la.item() - lb.item()) <= 1e-3 * abs(la.item())     # (the built-in means are taken over the bf16 rows)
    assert _cos(xa.grad, xb.grad) >= 0.9999


@pytest.mark.parametrize("n,d,c", [(65536, 128, 21), (4999, 64, 24), (131, 128, 3), (9000, 256, 21)])
def test_tensor_pipe_paths_through_the_c_abi(n, d, c):
    """bf16 rows take the tensor-pipe kernels (csrc/pool_tc.cuh).  Through the C ABI: the scratch variant of the forward
    (per-CTA parts, fixed-order sum) is bit-reproducible and agrees with the reduction variant and with fp64; the
    backward's fp32 de / dw (both, de only, dw only) agree with fp64 to the hi/lo split's accuracy."""
    from muscle_b200 import _lib
    from muscle_b200._call import call
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(n + d + c)
    e = F.normalize(torch.randn(n, d, generator=gen), dim=1).to(torch.bfloat16)
    w = torch.relu(torch.randn(n, c, generator=gen)) * torch.rand(n, c, generator=gen)
    dproto = torch.randn(c, d, generator=gen)
    eg, wg, dpg = e.to(dev), w.to(dev), dproto.to(dev)
    f32 = dict(dtype=torch.float32, device=dev)
    sb = int(lib.muscle_pool_prototypes_fwd_scratch_bytes(d, c))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    outs = []
    for _ in range(2):
        sums, mass, proto = torch.empty(c, d, **f32), torch.empty(c, **f32), torch.empty(c, d, **f32)
        scratch.fill_(0xff)
        call("muscle_pool_prototypes_fwd_ws", dev, eg.data_ptr(), 1, wg.data_ptr(), n, d, c, 1e-30, sums.data_ptr(),
             mass.data_ptr(), proto.data_ptr(), scratch.data_ptr(), sb)
        outs.append((sums.cpu(), mass.cpu(), proto.cpu()))
    if os.environ.get("MUSCLE_POOL_TC", "1") != "0":
        assert all(torch.equal(a, b) for a, b in zip(outs[0], outs[1]))       # fixed summation order
    sums2, mass2, proto2 = torch.empty(c, d, **f32), torch.empty(c, **f32), torch.empty(c, d, **f32)
    call("muscle_pool_prototypes_fwd", dev, eg.data_ptr(), 1, wg.data_ptr(), n, d, c, 1e-30, sums2.data_ptr(), mass2.data_ptr(),
         proto2.data_ptr())
    ed, wd = e.double(), w.double()
    ref_sum, ref_mass = wd.t() @ ed, wd.sum(0)
    scale = float(ref_sum.abs().max())
    for got_s, got_m in ((outs[0][0], outs[0][1]), (sums2.cpu(), mass2.cpu())):
        torch.testing.assert_close(got_s.double(), ref_sum, rtol=1e-4, atol=2e-5 * scale)
        torch.testing.assert_close(got_m.double(), ref_mass, rtol=1e-5, atol=0)
    proto, mass = outs[0][2].to(dev), outs[0][1].to(dev)
    g = dproto.double() / ref_mass[:, None]
    ref_de = wd @ g
    ref_dw = ed @ g.t() - (g * (ref_sum / ref_mass[:, None])).sum(1)[None, :]
    for want_de, want_dw in ((True, True), (True, False), (False, True)):
        de = torch.full((n, d), float("nan"), **f32) if want_de else None
        dw = torch.full((n, c), float("nan"), **f32) if want_dw else None
        call("muscle_pool_prototypes_bwd", dev, eg.data_ptr(), 1, wg.data_ptr(), n, d, c, 1e-30, proto.data_ptr(),
             mass.data_ptr(), dpg.data_ptr(), None if de is None else de.data_ptr(), None if dw is None else dw.data_ptr())
        if want_de:
            torch.testing.assert_close(de.cpu().double(), ref_de, rtol=1e-3, atol=1e-5 * float(ref_de.abs().max()))
        if want_dw:
            torch.testing.assert_close(dw.cpu().double(), ref_dw, rtol=1e-3, atol=2e-5 * float(ref_dw.abs().max()))


def test_tensor_pipe_forward_with_unaligned_weights():
    """The tensor-pipe forward stages a chunk's weights with one bulk copy when `w` is 16-byte aligned; a `w` that is not
    (here: a row-offset view, 84 bytes in) is loaded directly.  Both give the same sums."""
    from muscle_b200 import _lib
    from muscle_b200._call import call
    _lib.load()
    dev = torch.device("cuda", 0)
    n, d, c = 7001, 128, 21
    gen = torch.Generator().manual_seed(5)
    e = F.normalize(torch.randn(n, d, generator=gen), dim=1).to(torch.bfloat16).to(dev)
    w_big = torch.rand(n + 1, c, generator=gen).to(dev)
    w_off = w_big[1:]                                   # data_ptr 84 bytes past an aligned address
    w_al = w_off.clone()
    assert w_off.data_ptr() % 16 != 0 and w_al.data_ptr() % 16 == 0 and w_off.is_contiguous()
    f32 = dict(dtype=torch.float32, device=dev)
    res = []
    for w in (w_off, w_al):
        sums, mass, proto = torch.empty(c, d, **f32), torch.empty(c, **f32), torch.empty(c, d, **f32)
        call("muscle_pool_prototypes_fwd", dev, e.data_ptr(), 1, w.data_ptr(), n, d, c, 1e-30, sums.data_ptr(), mass.data_ptr(),
             proto.data_ptr())
        res.append((sums.cpu(), mass.cpu()))
    ref = w_al.double().t().cpu() @ e.double().cpu()
    for s_, m_ in res:
        torch.testing.assert_close(s_.double(), ref, rtol=1e-4, atol=2e-5 * float(ref.abs().max()))
        torch.testing.assert_close(m_.double(), w_al.double().sum(0).cpu(), rtol=1e-5, atol=0)
