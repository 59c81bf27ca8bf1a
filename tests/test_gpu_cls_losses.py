"""f4: FocalLoss, Log_Sum_Exp_Pairwise_Loss and the soft-margin term (src/loss_multilabel.py:24-33,68-91; composed
at train_mcl.py:180-183) against the reference-minted fixture and the oracle."""
import pytest
import torch

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _close(a, b, rtol=1e-4, atol=1e-6):
    torch.testing.assert_close(a.detach().cpu().float(), b.detach().cpu().float(), rtol=rtol, atol=atol)


def test_cls_losses_golden(golden):
    from muscle_b200 import loss_multilabel as L
    g = golden("cls_losses")
    logits = g["logits"].cuda().requires_grad_()
    loss, lf, ls, lp = L.mcl_cls_loss(logits, g["label"].cuda())
    _close(lf, g["loss_focal"])
    _close(ls, g["loss_softmargin"])
    _close(lp, g["loss_pair"])
    _close(loss, g["loss_cls"])
    (3.0 * loss).backward()
    _close(logits.grad, 3.0 * g["g_logits"], rtol=1e-3, atol=1e-6)
    assert float(logits.grad[:, 0].abs().max()) == 0.0            # the background logit takes no part
    # the two reference callables on their own inputs, composed by torch exactly as the trainer does
    lg = g["logits"].cuda().requires_grad_()
    prob = torch.sigmoid(lg[:, 1:])
    focal = L.FocalLoss()(prob, g["label"].cuda())
    rows = L.Log_Sum_Exp_Pairwise_Loss(prob, g["label"].cuda())
    _close(focal, g["loss_focal"])
    _close(rows, g["pair_rows"])
    soft = torch.nn.MultiLabelSoftMarginLoss()(lg[:, 1:], g["label"].cuda())
    (rows.mean() + soft + focal).backward()
    _close(lg.grad, g["g_logits"], rtol=1e-3, atol=1e-6)


@pytest.mark.parametrize("batch,c,with_bg", [(16, 20, True), (64, 20, True), (7, 5, False), (256, 20, True), (1, 32, False)])
def test_cls_losses_oracle(batch, c, with_bg):
    from muscle_b200 import loss_multilabel as L
    gen = torch.Generator().manual_seed(batch + c)
    logits = torch.randn(batch, c + (1 if with_bg else 0), generator=gen) * 3
    label = (torch.rand(batch, c, generator=gen) < 0.2).float()
    lr = logits.clone().requires_grad_()
    fg = lr[:, 1:] if with_bg else lr
    prob = torch.sigmoid(fg)
    ref = O.lsep_loss(prob, label).mean() + O.soft_margin_loss(fg, label) + O.focal_loss(prob, label)
    ref.backward()
    lg = logits.cuda().requires_grad_()
    loss, lf, ls, lp = L.mcl_cls_loss(lg, label.cuda())
    loss.backward()
    _close(loss, ref, rtol=2e-4)
    _close(lg.grad, lr.grad, rtol=2e-3, atol=1e-6)
