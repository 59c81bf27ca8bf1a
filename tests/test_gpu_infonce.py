"""GPU parity tests of the dense InfoNCE path (through the C ABI) against the CPU oracle and the
fixtures minted by the unmodified reference.  Tolerances are north_star's: loss within 1e-3
relative of the fp32 reference, input-gradient cosine >= 0.999."""
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-3
GRAD_COS = 0.999


@pytest.fixture(scope="module")
def M():
    import muscle_b200
    return muscle_b200


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


@pytest.mark.parametrize("d", [64, 128, 256])
def test_tcgen05_tile_blocks(M, d):
    torch.manual_seed(d)
    a = (torch.randn(128, d) / d ** 0.5).to(torch.bfloat16)
    b = (torch.randn(128, d) / d ** 0.5).to(torch.bfloat16)
    from muscle_b200.ops import debug_tile       # libmuscle_b200_diag.so: a test fixture, not in the product library
    s, g = debug_tile(a.cuda(), b.cuda())
    s_ref = a.float() @ b.float().t()
    g_ref = s_ref.to(torch.bfloat16).float() @ b.float()
    torch.testing.assert_close(s.cpu(), s_ref, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(g.cpu(), g_ref, rtol=1e-2, atol=1e-2)


def _run_both(M, x, cls, num_classes=21, temperature=0.1, nchw=True):
    """x [B,D,h,w] cpu fp32, cls [B,h,w]."""
    d = x.shape[1]
    xr = x.clone().requires_grad_()
    ref = O.dense_infonce_from_features(xr.permute(0, 2, 3, 1).reshape(-1, d), cls.reshape(-1),
                                        temperature, num_classes)
    gref, = torch.autograd.grad(ref, xr)
    if nchw:
        xg = x.cuda().requires_grad_()
        loss = M.dense_infonce(xg, cls.cuda(), temperature, num_classes)
        loss.backward()
        g = xg.grad.cpu()
    else:
        xg = x.permute(0, 2, 3, 1).reshape(-1, d).contiguous().cuda().requires_grad_()
        loss = M.dense_infonce(xg, cls.reshape(-1).cuda(), temperature, num_classes)
        loss.backward()
        g = xg.grad.cpu().reshape(x.shape[0], x.shape[2], x.shape[3], d).permute(0, 3, 1, 2)
    return loss.item(), ref.item(), g, gref


@pytest.mark.parametrize("batch,d,h,w,nchw", [
    (2, 128, 16, 16, True), (2, 128, 16, 16, False), (2, 128, 32, 32, True),   # config 1 = (2,128,32,32)
    (3, 64, 20, 24, True), (2, 256, 32, 32, True), (1, 128, 7, 9, True), (2, 128, 19, 23, False),
])
def test_dense_infonce_matches_oracle(M, batch, d, h, w, nchw):
    x, cls, _ = O.synth_embeddings(batch, d, h, w, seed=100 + batch + d + h)
    loss, ref, g, gref = _run_both(M, x, cls, nchw=nchw)
    assert abs(loss - ref) <= LOSS_RTOL * abs(ref), (loss, ref)
    assert _cos(g, gref) >= GRAD_COS
    assert abs(g.norm().item() / gref.norm().item() - 1) < 2e-2


@pytest.mark.parametrize("d", [64, 128])
def test_dense_infonce_golden_fixture(M, golden, d):
    """The CUDA path against the REFERENCE itself: the fixture is the reference info_nce called once per class
    (oracle/make_golden.py:gen_dense_infonce), run in fp32 ("the reference fp32 path") and in fp64."""
    g = golden("dense_infonce")["d%d" % d]
    xg = g["x"].detach().cuda().requires_grad_()
    loss = M.dense_infonce(xg, g["cls"].cuda(), 0.1, g["num_classes"])
    loss.backward()
    for tag in ("f32", "f64"):
        ref = g["loss_" + tag].item()
        assert abs(loss.item() - ref) <= LOSS_RTOL * abs(ref), (tag, loss.item(), ref)
        assert _cos(xg.grad.cpu(), g["gx_" + tag]) >= GRAD_COS
    assert abs(xg.grad.cpu().norm().item() / g["gx_f64"].norm().item() - 1) < 1e-2


@pytest.mark.parametrize("temperature", [0.02, 0.05, 0.07, 0.5])
def test_temperature_range(M, temperature):
    """The label mask as an exponent shift must not leak at small temperatures (the shift follows the
    temperature, csrc/pair_kernel.cuh:pair_mask_shift).  All embeddings share a common direction, so every
    similarity is high (~0.75 across classes, ~0.9 within): masked columns are the largest terms there are, a
    leak would dominate the row sums, and the loss stays O(1) at every temperature."""
    gen = torch.Generator().manual_seed(5)
    b, d, h, w = 2, 128, 24, 24
    _, cls, _ = O.synth_embeddings(b, d, h, w, seed=41)
    centers = torch.randn(21, d, generator=gen)
    base = torch.randn(d, generator=gen)
    x = (1.7 * base[None, :, None, None] + 0.7 * centers[cls].permute(0, 3, 1, 2)
         + 0.45 * torch.randn(b, d, h, w, generator=gen))
    loss, ref, g, gref = _run_both(M, x, cls, temperature=temperature)
    assert ref > 1e-2, ref
    assert abs(loss - ref) <= LOSS_RTOL * abs(ref), (loss, ref)
    assert _cos(g, gref) >= GRAD_COS


def test_temperature_below_the_fp32_range_is_refused(M):
    x, cls, _ = O.synth_embeddings(1, 128, 8, 8, seed=3)
    with pytest.raises(RuntimeError, match="temperature"):
        M.dense_infonce(x.cuda(), cls.cuda(), 0.005, 21)


def test_ignore_pixels_and_single_class(M):
    x, cls, _ = O.synth_embeddings(2, 128, 16, 16, seed=7)
    cls = cls.clone()
    cls[0, :4] = -1                      # ignored pixels: neither queries nor keys
    loss, ref, g, gref = _run_both(M, x, cls)
    assert abs(loss - ref) <= LOSS_RTOL * abs(ref)
    assert _cos(g, gref) >= GRAD_COS
    assert float(g[0, :, :4].abs().max()) == 0.0
    # one class only: no negatives -> every query's loss is exactly 0 (info_nce with K=0)
    one = torch.zeros_like(cls)
    xg = x.cuda().requires_grad_()
    l1 = M.dense_infonce(xg, one.cuda(), 0.1, 21)
    l1.backward()
    assert abs(l1.item()) < 1e-5
    assert float(xg.grad.abs().max()) < 1e-6


def test_forward_only_and_determinism_of_value(M):
    x, cls, _ = O.synth_embeddings(2, 128, 24, 24, seed=11)
    a = M.dense_infonce(x.cuda(), cls.cuda(), 0.1, 21)
    b = M.dense_infonce(x.cuda(), cls.cuda(), 0.1, 21)
    assert not a.requires_grad
    assert abs(a.item() - b.item()) <= 1e-5 * abs(a.item())


def test_upstream_gradient_is_applied_in_backward(M):
    """The gradient pass runs in backward() with autograd's upstream gradient multiplied in by the
    last kernel (muscle_dense_infonce_bwd's grad_scale): d(2.5 L)/dx == 2.5 dL/dx, both layouts."""
    x, cls, _ = O.synth_embeddings(2, 128, 20, 20, seed=17)
    for to_layout in (lambda t: t, lambda t: t.permute(0, 2, 3, 1).reshape(-1, 128).contiguous()):
        xa = to_layout(x).cuda().requires_grad_()
        xb = to_layout(x).cuda().requires_grad_()
        c = cls.reshape(-1).cuda() if xa.dim() == 2 else cls.cuda()
        M.dense_infonce(xa, c, 0.1, 21).backward()
        (2.5 * M.dense_infonce(xb, c, 0.1, 21)).backward()
        # two separate forward passes: the row sums are fp32 atomics (order varies from run to run), so the
        # gradients agree to fp32 summation noise, not bit for bit
        assert torch.allclose(xb.grad, 2.5 * xa.grad, rtol=1e-3, atol=1e-8)
        assert _cos(xb.grad, 2.5 * xa.grad) >= 0.9999999
    # the fused C entry point and the fwd / bwd pair agree (row sums are float atomics: not bit for bit)
    import muscle_b200._lib as L
    lib = L.load()
    xc = x.cuda().contiguous()
    n, dim = 2 * 20 * 20, 128
    wsb = lib.muscle_dense_infonce_workspace_bytes(n, dim, 21)
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    loss = torch.empty((), device="cuda")
    dx = torch.empty_like(xc)
    sp = torch.cuda.current_stream().cuda_stream
    cc = cls.cuda().contiguous()
    rc = lib.muscle_dense_infonce_fwd_bwd(xc.data_ptr(), 1, 2, 400, dim, cc.data_ptr(), 1, 21, 0.1, 1e-12,
                                          loss.data_ptr(), dx.data_ptr(), ws.data_ptr(), wsb, sp)
    assert rc == 0
    xg = xc.clone().requires_grad_()
    l2 = M.dense_infonce(xg, cc, 0.1, 21)
    l2.backward()
    assert abs(l2.item() - loss.item()) <= 1e-5 * abs(loss.item())
    assert torch.allclose(xg.grad, dx, rtol=1e-3, atol=1e-8)


def test_value_and_grad_matches_autograd_path(M):
    """dense_infonce_value_and_grad (one fused C call, no autograd) == dense_infonce + backward, both
    layouts, and against the oracle."""
    x, cls, _ = O.synth_embeddings(2, 128, 24, 20, seed=23)
    xr = x.clone().requires_grad_()
    ref = O.dense_infonce_from_features(xr.permute(0, 2, 3, 1).reshape(-1, 128), cls.reshape(-1), 0.1, 21)
    gref, = torch.autograd.grad(ref, xr)
    for to_layout in (lambda t: t, lambda t: t.permute(0, 2, 3, 1).reshape(-1, 128).contiguous()):
        xa = to_layout(x).cuda().requires_grad_()
        c = cls.reshape(-1).cuda() if xa.dim() == 2 else cls.cuda()
        la = M.dense_infonce(xa, c, 0.1, 21)
        la.backward()
        lv, dx = M.dense_infonce_value_and_grad(to_layout(x).cuda(), c, 0.1, 21)
        assert dx.shape == xa.shape and not lv.requires_grad
        assert abs(lv.item() - la.item()) <= 1e-5 * abs(la.item())
        assert torch.allclose(dx, xa.grad, rtol=1e-3, atol=1e-8)
        assert abs(lv.item() - ref.item()) <= LOSS_RTOL * abs(ref.item())
        assert _cos(dx.cpu(), to_layout(gref)) >= GRAD_COS
    with pytest.raises(RuntimeError):
        M.dense_infonce_value_and_grad(torch.randn(64, 128), torch.zeros(64, dtype=torch.int64))


def test_permutation_invariance_and_temperature(M):
    x, cls, _ = O.synth_embeddings(2, 128, 16, 16, seed=13)
    flat = x.permute(0, 2, 3, 1).reshape(-1, 128)
    c = cls.reshape(-1)
    perm = torch.randperm(flat.shape[0], generator=torch.Generator().manual_seed(1))
    l0 = M.dense_infonce(flat.cuda(), c.cuda(), 0.07, 21).item()
    l1 = M.dense_infonce(flat[perm].contiguous().cuda(), c[perm].cuda(), 0.07, 21).item()
    ref = O.dense_infonce_from_features(flat, c, 0.07, 21).item()
    assert abs(l0 - l1) <= 1e-4 * abs(l0)
    assert abs(l0 - ref) <= LOSS_RTOL * abs(ref)


@pytest.mark.parametrize("bg_bias", [0.0, 0.15])
def test_config2_full_size_value_and_gradient(M, bg_bias):
    """BASELINE config 2 (B=8, 448 crop -> 56x56, D=128) as bench.py generates it (bg_bias 0) and as round 1
    did (0.15, ~75 % background): value AND input gradient against the row-chunked oracle (north_star: 1e-3
    relative, cosine >= 0.999), gradient orthogonal to x (normalisation), finite, and equal across layouts."""
    import workloads
    x, cls, _ = workloads.synth_embeddings(8, 128, 56, 56, seed=1234, bg_bias=bg_bias)
    xg = x.cuda().requires_grad_()
    loss = M.dense_infonce(xg, cls.cuda(), 0.1, 21)
    loss.backward()
    g = xg.grad
    assert torch.isfinite(g).all()
    radial = (g * xg.detach()).sum(1).abs().max().item()
    assert radial <= 1e-3 * g.norm(dim=1).max().item() * xg.detach().norm(dim=1).max().item()
    flat = x.permute(0, 2, 3, 1).reshape(-1, 128)
    ref, gref = O.dense_infonce_value_and_grad_chunked(flat, cls.reshape(-1), 0.1, 21)
    assert abs(loss.item() - ref) <= LOSS_RTOL * abs(ref)
    g_flat = g.permute(0, 2, 3, 1).reshape(-1, 128).cpu()
    assert _cos(g_flat, gref) >= GRAD_COS
    assert abs(g_flat.norm().item() / gref.norm().item() - 1) < 1e-2
    # per pixel, not only in aggregate: 99 % of the pixels' own gradient vectors point the reference's way
    pix_cos = F.cosine_similarity(g_flat.double(), gref.double(), dim=1)
    assert float((pix_cos >= 0.99).float().mean()) >= 0.99
    xp = flat.contiguous().cuda().requires_grad_()
    lp = M.dense_infonce(xp, cls.reshape(-1).cuda(), 0.1, 21)
    lp.backward()
    gp = xp.grad.reshape(8, 56, 56, 128).permute(0, 3, 1, 2)
    assert abs(lp.item() - loss.item()) <= 1e-5 * abs(loss.item())
    assert _cos(gp.cpu(), g.cpu()) >= 0.99999


def test_config3_full_size_value_gradient_and_relabelling(M):
    """BASELINE config 3 (B=16, 512 crop -> 64x64, D=128; 65 536 pixels, 4.3e9 pairs): value and input gradient
    against the row-chunked oracle (forward and backward); gradient finite and orthogonal to x; and a size-independent
    property of the loss — renaming the classes (a permutation of the 21 labels, which reorders
    the class-sorted layout and every tile's mask) changes neither the value nor the gradient."""
    x, cls, _ = O.synth_embeddings(16, 128, 64, 64, seed=77)
    xg = x.cuda().requires_grad_()
    loss = M.dense_infonce(xg, cls.cuda(), 0.1, 21)
    loss.backward()
    g = xg.grad
    assert torch.isfinite(g).all()
    radial = (g * xg.detach()).sum(1).abs().max().item()
    assert radial <= 1e-3 * g.norm(dim=1).max().item() * xg.detach().norm(dim=1).max().item()
    ref, gref = O.dense_infonce_value_and_grad_chunked(x.permute(0, 2, 3, 1).reshape(-1, 128), cls.reshape(-1), 0.1, 21,
                                                       row_chunk=4096)
    assert abs(loss.item() - ref) <= LOSS_RTOL * abs(ref), (loss.item(), ref)
    g_flat = g.permute(0, 2, 3, 1).reshape(-1, 128).cpu()
    assert _cos(g_flat, gref) >= GRAD_COS               # all 65 536 pixels against the chunked oracle backward
    assert abs(g_flat.norm().item() / gref.norm().item() - 1) < 1e-2
    pi = torch.randperm(21, generator=torch.Generator().manual_seed(3))
    relabelled = torch.where(cls >= 0, pi[cls.clamp(min=0)], cls)
    xr = x.cuda().requires_grad_()
    lr = M.dense_infonce(xr, relabelled.cuda(), 0.1, 21)
    lr.backward()
    assert abs(lr.item() - loss.item()) <= 1e-5 * abs(loss.item()), (lr.item(), loss.item())
    assert _cos(xr.grad.cpu(), g.cpu()) >= 0.99999


def test_errors_are_loud(M):
    x = torch.randn(64, 96, device="cuda")
    with pytest.raises((RuntimeError, ValueError)):
        M.dense_infonce(x, torch.zeros(64, dtype=torch.int64, device="cuda"))
    with pytest.raises(RuntimeError):
        M.dense_infonce(torch.randn(64, 128), torch.zeros(64, dtype=torch.int64))
