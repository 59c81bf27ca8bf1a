"""IRN random walk (src/indexing.py:116-150) against the reference-minted fixture."""
import pytest
import torch

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def test_transition_matrix_golden(golden):
    from muscle_b200 import indexing
    g = golden("irn")
    a = g["aff"].cuda()
    for times in (1, 4, 8):
        t = indexing.to_transition_matrix(a, 10, times)
        torch.testing.assert_close(t.cpu(), g["trans_%d" % times], rtol=2e-4, atol=1e-9)
    torch.testing.assert_close(indexing.to_transition_matrix(a, 2, 3).cpu(), g["trans_b2_3"], rtol=2e-4, atol=1e-9)
    assert abs(float(indexing.to_transition_matrix(a, 10, 0).sum(0).mean()) - 1.0) < 1e-5      # columns sum to 1


def test_random_walk_matches_oracle():
    from muscle_b200 import indexing
    gen = torch.Generator().manual_seed(3)
    h, w = 24, 20
    a = torch.rand(h * w, h * w, generator=gen)
    aff = 0.5 * (a + a.t())
    aff.fill_diagonal_(1.0)
    x = torch.rand(5, h, w, generator=gen)
    edge = torch.rand(h, w, generator=gen)
    t_ref = O.to_transition_matrix(aff, 10, 6)
    rw_ref = O.random_walk(x, edge, t_ref)
    t = indexing.to_transition_matrix(aff.cuda(), 10, 6)
    rw = indexing.random_walk(x.cuda(), edge.cuda(), t)
    torch.testing.assert_close(t.cpu(), t_ref, rtol=2e-4, atol=1e-9)
    torch.testing.assert_close(rw.cpu(), rw_ref, rtol=2e-4, atol=1e-7)
