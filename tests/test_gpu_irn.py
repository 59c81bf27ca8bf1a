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


@pytest.mark.parametrize("m,k,n", [(300, 300, 300), (37, 1000, 1000), (128, 256, 384), (5, 130, 70)])
def test_split3_products_are_fp32_accurate(m, k, n):
    """The squarings run as three-part bf16 products on the tensor pipe (csrc/irn.cu): against fp64 on non-negative
    operands the result is as accurate as an fp32 GEMM (no cancellation: relative error per entry ~1e-6), both as the
    fp32 output and as the re-split parts handed to the next squaring."""
    from muscle_b200 import _lib
    from muscle_b200._call import call
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(m + n)
    a = torch.rand(m, k, generator=gen) ** 6            # wide dynamic range, like A ** beta
    b = torch.rand(k, n, generator=gen) ** 6
    pad = lambda v: (v + 127) // 128 * 128
    mp, kp, np_ = pad(m), pad(k), pad(n)
    ag, bg = a.to(dev), b.to(dev)
    pa = torch.empty(int(lib.muscle_irn_split3_bytes(mp, kp)), dtype=torch.uint8, device=dev)
    pb = torch.empty(int(lib.muscle_irn_split3_bytes(kp, np_)), dtype=torch.uint8, device=dev)
    pc = torch.full((int(lib.muscle_irn_split3_bytes(mp, np_)),), 0xff, dtype=torch.uint8, device=dev)
    call("muscle_irn_split3", dev, ag.data_ptr(), m, k, mp, kp, pa.data_ptr())
    call("muscle_irn_split3", dev, bg.data_ptr(), k, n, kp, np_, pb.data_ptr())
    c = torch.full((m, n), float("nan"), device=dev)
    call("muscle_irn_matmul_split3", dev, pa.data_ptr(), pb.data_ptr(), mp, kp, np_, m, n, c.data_ptr(), pc.data_ptr())
    ref = a.double() @ b.double()
    torch.testing.assert_close(c.cpu().double(), ref, rtol=1e-5, atol=0)
    parts = pc.view(torch.bfloat16).reshape(3, mp, np_).float()
    again = parts.sum(0)                                   # h + m + l reproduces the fp32 result
    torch.testing.assert_close(again[:m, :n].cpu(), c.cpu(), rtol=2e-7, atol=1e-30)
    assert float(again[m:].abs().max() if mp > m else 0.0) == 0.0 and float(again[:, n:].abs().max() if np_ > n else 0.0) == 0.0


def test_library_gemm_switch_agrees(monkeypatch, golden):
    from muscle_b200 import indexing
    g = golden("irn")
    a = g["aff"].cuda()
    t_tc = indexing.to_transition_matrix(a, 10, 8)
    monkeypatch.setenv("MUSCLE_IRN_TC", "0")
    t_lib = indexing.to_transition_matrix(a, 10, 8)
    torch.testing.assert_close(t_tc, t_lib, rtol=5e-5, atol=1e-12)
