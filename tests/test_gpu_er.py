"""SURVEY.md §8f rank 1 on the GPU: bilinear up-sampling (align_corners=True; src/MuSCLe.py:256-257) forward and
adjoint, and the ER loss of train_mcl.py:175-188 (radix-select top-k, no host sync) against the reference-minted
fixture and the oracle."""
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


@pytest.mark.parametrize("shape,size", [((2, 21, 7, 9), (28, 36)), ((3, 5, 56, 56), (448, 448)), ((1, 4, 16, 16), (16, 16)),
                                        ((2, 3, 40, 30), (20, 15)), ((2, 2, 1, 5), (7, 9)), ((1, 256, 16, 16), (128, 128))])
def test_upsample_forward_and_adjoint(shape, size):
    from muscle_b200.er import upsample_bilinear
    gen = torch.Generator().manual_seed(shape[2] * 7 + size[0])
    x = torch.randn(*shape, generator=gen)
    go = torch.randn(*shape[:2], *size, generator=gen)
    xr = x.clone().requires_grad_()
    yr = F.interpolate(xr, size=size, mode="bilinear", align_corners=True)
    (yr * go).sum().backward()
    xg = x.cuda().requires_grad_()
    y = upsample_bilinear(xg, size)
    (y * go.cuda()).sum().backward()
    torch.testing.assert_close(y.detach().cpu(), yr.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(xg.grad.cpu(), xr.grad, rtol=1e-4, atol=1e-5)


def test_er_loss_golden(golden):
    from muscle_b200.er import er_loss, upsample_bilinear
    g = golden("er_loss")
    sg = g["sgc_small"].cuda().requires_grad_()
    up_c = upsample_bilinear(g["cams_small"].cuda(), g["size"])
    up_s = upsample_bilinear(sg, g["size"])
    torch.testing.assert_close(up_c.cpu(), g["up_cams"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(up_s.detach().cpu(), g["up_sgc"], rtol=1e-5, atol=1e-6)
    loss = er_loss(up_c, up_s, g["label"].cuda())
    loss.backward()
    assert abs(loss.item() - g["loss_er"].item()) <= 1e-4 * abs(g["loss_er"].item()), (loss.item(), g["loss_er"].item())
    assert _cos(sg.grad.cpu(), g["g_sgc_small"]) >= 0.9999
    torch.testing.assert_close(sg.grad.cpu(), g["g_sgc_small"], rtol=1e-3, atol=1e-3 * float(g["g_sgc_small"].abs().max()))


@pytest.mark.parametrize("n,h,w", [(4, 64, 64), (2, 37, 53), (8, 224, 224)])
def test_er_loss_oracle(n, h, w):
    from muscle_b200.er import er_loss
    gen = torch.Generator().manual_seed(n * h)
    label = O.synth_labels(n, gen)
    cams = F.avg_pool2d(torch.randn(n, 21, h, w, generator=gen), 5, 1, 2) * 6
    sgcs = cams + 1.5 * F.avg_pool2d(torch.randn(n, 21, h, w, generator=gen), 5, 1, 2)
    sr = sgcs.clone().requires_grad_()
    ref = O.er_loss(cams, sr, label)
    ref.backward()
    sg = sgcs.cuda().requires_grad_()
    loss = er_loss(cams.cuda(), sg, label.cuda())
    (2.0 * loss).backward()
    assert abs(loss.item() - ref.item()) <= 1e-4 * abs(ref.item()), (loss.item(), ref.item())
    assert _cos(sg.grad.cpu(), 2 * sr.grad) >= 0.9999
    # an image without any label (valid_channel = 0 for the whole batch) -> k = 0 -> NaN like torch.topk(k=0).mean()
    z = er_loss(cams[:1].cuda(), sgcs[:1].cuda(), torch.zeros(1, 20).cuda())
    assert z.item() != z.item()
