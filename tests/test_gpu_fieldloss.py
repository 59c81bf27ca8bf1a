"""GPU parity of FieldLoss against the reference-minted fixture and
the CPU oracle, consuming Python's `random` stream exactly like the reference."""
import random

import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def test_fieldloss_stages_golden(golden):
    from muscle_b200.edge import FieldLoss
    g = golden("fieldloss")
    crit = FieldLoss(sobel_size=5, beta=g["beta"], k=g["k"])
    st = crit.edges(g["seg"].cuda(), g["label_bg"].cuda(), g["step"])
    torch.testing.assert_close(st["mag_sum"].cpu(), g["mag"], rtol=1e-4, atol=1e-4)
    present = g["label_bg"][:, 1:].bool()[:, :, None, None].expand_as(g["orient"])
    agree = (st["orient"].cpu().float() == g["orient"])[present].float().mean().item()
    assert agree >= 0.999, agree
    assert torch.equal(st["pos"].cpu().bool(), g["pos"].bool())          # index work: exact
    # in/out lists of (image 0, class 14): run the full forward to materialise them
    random.seed(g["py_seed"])
    crit(g["seg"].cuda(), g["dense"].cuda(), g["mask"].cuda(), g["label_bg"].cuda(), g["step"])
    last = crit._last
    i = last["planes"].index(0 * 20 + 14)
    lists = last["lists"].cpu()
    o0, o1 = last["offsets"][2 * i], last["offsets"][2 * i + 1]
    n_in = g["in_idx_0_14"].numel()
    assert torch.equal(lists[o0:o1].long(), g["out_idx_0_14"].long())
    assert torch.equal(lists[o1:o1 + n_in].long(), g["in_idx_0_14"].long())


def test_fieldloss_golden_loss_and_grad(golden):
    from muscle_b200.edge import FieldLoss
    g = golden("fieldloss")
    crit = FieldLoss(sobel_size=5, beta=g["beta"], k=g["k"])
    dense = g["dense"].cuda().requires_grad_()
    random.seed(g["py_seed"])
    loss, mag = crit(g["seg"].cuda(), dense, g["mask"].cuda(), g["label_bg"].cuda(), g["step"])
    assert torch.is_tensor(loss) and loss.dim() == 0
    loss.backward()
    assert abs(loss.item() - g["loss"].item()) <= 1e-3 * abs(g["loss"].item()), (loss.item(), g["loss"].item())
    assert _cos(dense.grad.cpu(), g["gd"]) >= 0.999
    torch.testing.assert_close(dense.grad.cpu(), g["gd"], rtol=1e-2, atol=1e-3 * float(g["gd"].abs().max()))
    torch.testing.assert_close(mag.cpu(), g["mag"], rtol=1e-4, atol=1e-4)


def test_fieldloss_gates(golden):
    from muscle_b200.edge import FieldLoss
    g = golden("fieldloss")
    gen = torch.Generator().manual_seed(1)
    crit = FieldLoss(sobel_size=5, beta=1e2, k=g["k"])
    flat = torch.zeros(1, 21, 24, 24).cuda()
    only_bg = torch.cat([torch.ones(1, 1), torch.zeros(1, 20)], 1).cuda()
    loss0, mag0 = crit(flat, torch.randn(1, 24, 24, 24, generator=gen).cuda(), flat, only_bg, g["step"])
    assert loss0 is False and g["gate_false"]
    assert mag0.shape == (1, 24, 24)
    crit3 = FieldLoss(sobel_size=5, beta=1e2, k=100000)
    loss1, _ = crit3(g["seg"].cuda(), g["dense"].cuda(), g["mask"].cuda(), g["label_bg"].cuda(), g["step"])
    assert (not torch.is_tensor(loss1)) and loss1 == 0 and g["big_k_is_zero"]
    with pytest.raises(RuntimeError):
        crit(g["seg"], g["dense"], g["mask"], g["label_bg"], g["step"])      # CPU tensors: no fallback


@pytest.mark.parametrize("n,h,w,ch,k,step", [(2, 64, 64, 256, 16, 3), (3, 96, 80, 64, 24, 5)])
def test_fieldloss_matches_oracle(n, h, w, ch, k, step):
    from muscle_b200.edge import FieldLoss
    gen = torch.Generator().manual_seed(h + ch)
    label = O.synth_labels(n, gen)
    label_bg = torch.cat([torch.ones(n, 1), label], 1)
    noise = F.avg_pool2d(torch.randn(n, 21, h, w, generator=gen), 9, 1, 4, count_include_pad=False)
    noise = noise.masked_fill(~label_bg.bool()[:, :, None, None], -1e4)
    seg = noise * 3 + 0.05 * torch.randn(n, 21, h, w, generator=gen)
    mask = noise * 2 + 0.3 * torch.randn(n, 21, h, w, generator=gen)
    dense = torch.randn(n, ch, h, w, generator=gen)
    dr = dense.clone().requires_grad_()
    random.seed(5)
    ref, mag_r = O.fieldloss_forward(seg, dr, mask, label_bg, step=step, k=k, beta=1e2)
    dg = dense.cuda().requires_grad_()
    random.seed(5)
    out, mag = FieldLoss(sobel_size=5, beta=1e2, k=k)(seg.cuda(), dg, mask.cuda(), label_bg.cuda(), step)
    torch.testing.assert_close(mag.cpu(), mag_r, rtol=1e-4, atol=1e-4)
    assert torch.is_tensor(out) == torch.is_tensor(ref)
    if torch.is_tensor(ref):
        gref, = torch.autograd.grad(ref, dr)
        out.backward()
        assert abs(out.item() - ref.item()) <= 1e-3 * abs(ref.item()) + 1e-7, (out.item(), ref.item())
        assert _cos(dg.grad.cpu(), gref) >= 0.999
