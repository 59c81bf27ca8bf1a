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
    with pytest.raises(ValueError):
        info_nce(3 * qg.detach(), pg.detach(), kg.detach())
