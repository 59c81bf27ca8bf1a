"""GPU parity of the projection + PCM path against the reference-minted
fixture and the CPU oracle.  bf16 operands / fp32 accumulation: outputs within 1e-2 relative of
their scale and cosine >= 0.999 (north_star's gradient criterion) for every gradient."""
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def _check(name, got, ref, cos=0.999, rel=2e-2):
    got, ref = got.detach().cpu(), ref.detach().cpu()
    assert torch.isfinite(got).all(), name
    scale = float(ref.abs().max())
    err = float((got - ref).abs().max())
    print("%-8s cos %.6f  max|err|/max|ref| %.2e" % (name, _cos(got, ref), err / scale))
    assert _cos(got, ref) >= cos, (name, _cos(got, ref))
    assert err <= rel * scale, (name, err, scale)


def _run(cam, f, w, b, go):
    from muscle_b200 import heads
    cr, wr, br = cam.clone().requires_grad_(), w.clone().requires_grad_(), b.clone().requires_grad_()
    ref = O.pcm(cr, f, wr, br)
    (ref * go).sum().backward()
    cg, wg, bg = cam.cuda().requires_grad_(), w.cuda().requires_grad_(), b.cuda().requires_grad_()
    out = heads.pcm(cg, f.cuda(), wg, bg)
    (out * go.cuda()).sum().backward()
    return (out, ref), (cg.grad, cr.grad), (wg.grad, wr.grad), (bg.grad, br.grad)


def test_pcm_golden(golden):
    g = golden("heads")
    (out, _), (gc, _), (gw, _), (gb, _) = _run(g["pcm_cam"], g["pcm_f"], g["fuse_w"], g["fuse_b"], g["pcm_go"])
    _check("out", out, g["pcm_out"])
    _check("dcam", gc, g["pcm_gcam"])
    _check("dW", gw, g["pcm_gw"], rel=5e-2)
    _check("db", gb, g["pcm_gb"], rel=5e-2)


@pytest.mark.parametrize("n,cin,h,w", [(2, 208, 16, 16), (2, 336, 28, 28), (1, 64, 9, 13), (3, 208, 32, 32)])
def test_pcm_matches_oracle(n, cin, h, w):
    gen = torch.Generator().manual_seed(cin + h)
    cam = torch.rand(n, 21, h, w, generator=gen)
    f = torch.relu(torch.randn(n, cin, h, w, generator=gen))
    wt = torch.randn(128, cin, 1, 1, generator=gen) / cin ** 0.5
    bias = torch.randn(128, generator=gen) * 0.1
    go = torch.randn(n, 21, h, w, generator=gen)
    (out, ref), (gc, gcr), (gw, gwr), (gb, gbr) = _run(cam, f, wt, bias, go)
    _check("out", out, ref)
    _check("dcam", gc, gcr)
    _check("dW", gw, gwr, rel=5e-2)
    _check("db", gb, gbr, rel=5e-2)


def test_pcm_config2_stride8_properties():
    """BASELINE config 2 at stride 8 (B=8, 56x56, b7 Cin=336): rows of out are convex combinations
    of cam over pixels (S >= 0, normalised), so min cam <= out <= max cam per class; finite grads."""
    from muscle_b200 import heads
    gen = torch.Generator().manual_seed(8)
    cam = torch.rand(8, 21, 56, 56, generator=gen).cuda().requires_grad_()
    f = torch.relu(torch.randn(8, 336, 56, 56, generator=gen)).cuda()
    wt = (torch.randn(128, 336, 1, 1, generator=gen) / 336 ** 0.5).cuda().requires_grad_()
    bias = (torch.randn(128, generator=gen) * 0.1).cuda().requires_grad_()
    out = heads.pcm(cam, f, wt, bias)
    flat_c, flat_o = cam.detach().flatten(2), out.detach().flatten(2)
    assert (flat_o <= flat_c.max(-1, keepdim=True)[0] * 1.01 + 1e-3).all()
    assert (flat_o >= flat_c.min(-1, keepdim=True)[0] * 0.99 - 1e-3).all()
    out.square().mean().backward()
    for t in (cam.grad, wt.grad, bias.grad):
        assert torch.isfinite(t).all() and float(t.abs().max()) > 0
    # a subset of images against the oracle (the op is per image)
    ref = O.pcm(cam.detach()[:1].cpu(), f[:1].cpu(), wt.detach().cpu(), bias.detach().cpu())
    _check("out[0]", out[:1], ref)
