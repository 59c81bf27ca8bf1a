"""GPU parity of the InfoNCE primitive with distinct query / positive / negative sets (SURVEY.md §8a
row a1, src/loss_multilabel.py:9-21) against the reference-minted fixture and the oracle."""
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def test_info_nce_golden(golden):
    from muscle_b200.loss_multilabel import info_nce
    g = golden("info_nce")["f32"]
    q, p, k = (g[n].cuda().requires_grad_() for n in "qpk")
    loss = info_nce(q, p, k)
    loss.backward()
    assert abs(loss.item() - g["loss"].item()) <= 1e-3 * abs(g["loss"].item())
    for name, t in (("gq", q), ("gp", p), ("gk", k)):
        assert _cos(t.grad.cpu(), g[name]) >= 0.999, name
        assert abs(t.grad.norm().item() / g[name].norm().item() - 1) < 2e-2, name
    none = info_nce(q.detach(), p.detach(), k.detach(), temperature=0.07, reduction="none")
    torch.testing.assert_close(none.cpu(), g["none_t007"], rtol=5e-3, atol=5e-3)


@pytest.mark.parametrize("nq,npos,nk,d", [(300, 40, 1000, 128), (1000, 7, 513, 64), (257, 300, 2049, 256)])
def test_info_nce_matches_oracle(nq, npos, nk, d):
    from muscle_b200.loss_multilabel import info_nce
    gen = torch.Generator().manual_seed(nq + d)
    q = F.normalize(torch.randn(nq, d, generator=gen), dim=1)
    p = F.normalize(q[:1] + 0.5 * torch.randn(npos, d, generator=gen), dim=1)
    k = F.normalize(torch.randn(nk, d, generator=gen), dim=1)
    qr, pr, kr = (t.clone().requires_grad_() for t in (q, p, k))
    ref = O.info_nce(qr, pr, kr, reduction="sum")
    ref.backward()
    qg, pg, kg = (t.cuda().requires_grad_() for t in (q, p, k))
    out = info_nce(qg, pg, kg, reduction="sum")
    out.backward()
    assert abs(out.item() - ref.item()) <= 1e-3 * abs(ref.item())
    for a, b in ((qg, qr), (pg, pr), (kg, kr)):
        assert _cos(a.grad.cpu(), b.grad) >= 0.999
    with pytest.raises(RuntimeError):           # |q| |k| / T beyond the exponent range of the passes: refused, not wrong
        info_nce(30 * qg.detach(), pg.detach(), kg.detach())


@pytest.mark.parametrize("nq,npos,nk,d,tau", [(300, 40, 1000, 128, 0.5), (513, 9, 700, 64, 0.3), (200, 5, 600, 96, 1.0)])
def test_info_nce_rows_of_any_norm(nq, npos, nk, d, tau):
    """The reference never normalises (src/loss_multilabel.py:9-21): rows of any length, against the same expression
    evaluated by torch on the CPU (fp32)."""
    from muscle_b200.loss_multilabel import info_nce
    gen = torch.Generator().manual_seed(nq + d)
    q = torch.randn(nq, d, generator=gen) * (0.2 + torch.rand(nq, 1, generator=gen)) / d ** 0.5 * 1.5
    p = (q[:1] + 0.5 * torch.randn(npos, d, generator=gen) / d ** 0.5) * 0.8
    k = torch.randn(nk, d, generator=gen) * (0.1 + 1.2 * torch.rand(nk, 1, generator=gen)) / d ** 0.5
    qr, pr, kr = (t.clone().requires_grad_() for t in (q, p, k))
    ref = O.info_nce(qr, pr, kr, temperature=tau, reduction="mean")
    ref.backward()
    qg, pg, kg = (t.cuda().requires_grad_() for t in (q, p, k))
    out = info_nce(qg, pg, kg, temperature=tau, reduction="mean")
    out.backward()
    assert abs(out.item() - ref.item()) <= 1e-3 * abs(ref.item()), (out.item(), ref.item())
    for a, b in ((qg, qr), (pg, pr), (kg, kr)):
        assert _cos(a.grad.cpu(), b.grad) >= 0.999
        assert abs(a.grad.norm().item() / b.grad.norm().item() - 1) < 2e-2


@pytest.mark.parametrize("unit", [True, False])
def test_info_nce_reduction_none_backward(unit):
    """reduction='none' with an upstream gradient per row (F.cross_entropy(..., reduction='none') in the reference)"""
    from muscle_b200.loss_multilabel import info_nce
    gen = torch.Generator().manual_seed(11)
    nq, npos, nk, d, tau = 400, 12, 900, 128, 0.2
    q = torch.randn(nq, d, generator=gen)
    p = q[:1] + 0.5 * torch.randn(npos, d, generator=gen)
    k = torch.randn(nk, d, generator=gen)
    if unit:
        q, p, k = (F.normalize(t, dim=1) for t in (q, p, k))
    else:
        q, p, k = (t / d ** 0.5 * s for t, s in ((q, 0.9), (p, 0.7), (k, 1.1)))
    up = torch.randn(nq, generator=gen)
    qr, pr, kr = (t.clone().requires_grad_() for t in (q, p, k))
    rows_ref = O.info_nce(qr, pr, kr, temperature=tau, reduction="none")
    (rows_ref * up).sum().backward()
    qg, pg, kg = (t.cuda().requires_grad_() for t in (q, p, k))
    rows = info_nce(qg, pg, kg, temperature=tau, reduction="none")
    torch.testing.assert_close(rows.cpu(), rows_ref.detach(), rtol=5e-3, atol=5e-3)
    (rows * up.cuda()).sum().backward()
    for a, b in ((qg, qr), (pg, pr), (kg, kr)):
        assert _cos(a.grad.cpu(), b.grad) >= 0.999
        assert abs(a.grad.norm().item() / b.grad.norm().item() - 1) < 2e-2


def test_info_nce_any_norm_reference_fixture(golden):
    """rows of any norm and the backward of reduction='none' against outputs of the reference itself
    (tests/golden/info_nce_anynorm.pt, minted by oracle/make_golden.py from src/loss_multilabel.py:9-21)"""
    from muscle_b200.loss_multilabel import info_nce
    g = golden("info_nce_anynorm")["f32"]
    q, p, k = (g[n].cuda().requires_grad_() for n in "qpk")
    loss = info_nce(q, p, k, temperature=0.3)
    loss.backward()
    assert abs(loss.item() - g["loss"].item()) <= 1e-3 * abs(g["loss"].item())
    for name, t in (("gq", q), ("gp", p), ("gk", k)):
        assert _cos(t.grad.cpu(), g[name]) >= 0.999, name
        assert abs(t.grad.norm().item() / g[name].norm().item() - 1) < 2e-2, name
    q2, p2, k2 = (g[n].cuda().requires_grad_() for n in "qpk")
    rows = info_nce(q2, p2, k2, temperature=0.3, reduction="none")
    torch.testing.assert_close(rows.cpu(), g["rows"], rtol=5e-3, atol=5e-3)
    (rows * g["up"].cuda()).sum().backward()
    for name, t in (("nq", q2), ("np", p2), ("nk", k2)):
        assert _cos(t.grad.cpu(), g[name]) >= 0.999, name
        assert abs(t.grad.norm().item() / g[name].norm().item() - 1) < 2e-2, name
