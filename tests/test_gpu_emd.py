"""GPU parity of the region-matching path: crop feeder and Sinkhorn EMD
against the reference-minted fixture and the CPU oracle, with numpy's RNG consumed identically."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import muscle_oracle as O

pytestmark = pytest.mark.gpu


def _cos(a, b):
    return F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()


def _views(g):
    raw1 = g["raw1"].clone().requires_grad_()
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(g["raw2"] * 4, 1), dim=1)
    return raw1, vw1, vw2


def test_crops_match_reference_shapes_and_oracle_values(golden):
    from muscle_b200.torchutils import get_dynamic_crops
    g = golden("emd")
    _, vw1, vw2 = _views(g)
    np.random.seed(g["np_seed"])
    r1, r2, ridx = O.get_dynamic_crops(vw1.detach(), g["c1"], vw2, g["c2"])
    np.random.seed(g["np_seed"])
    c1, c2, idx = get_dynamic_crops(vw1.detach().cuda(), g["c1"], vw2.cuda(), g["c2"])
    assert idx == list(g["batch_indices"]) == ridx
    assert [[tuple(t.shape) for t in c] for c in c1] == g["crop_shapes1"]
    assert [[tuple(t.shape) for t in c] for c in c2] == g["crop_shapes2"]
    for ours, ref in ((c1, r1), (c2, r2)):
        for a, b in zip(ours, ref):
            for t, u in zip(a, b):
                torch.testing.assert_close(t.cpu(), u, rtol=1e-5, atol=1e-6)


def test_sinkhorn_value_golden(golden):
    """One isolated pair: the fixture's cost matrix comes from two crops; rebuild the regions from
    the same crops and compare the score with the reference's sinkhorn_logsumexp value."""
    from muscle_b200.loss_multilabel import EMD
    g = golden("emd")
    _, vw1, vw2 = _views(g)
    np.random.seed(g["np_seed"])
    r1, r2, _ = O.get_dynamic_crops(vw1.detach(), g["c1"], vw2, g["c2"])
    x, y = r1[0][0].cuda(), r2[0][1].cuda()
    val = EMD()([[x]], [[y]], mode="dynamic")
    assert abs(val.item() - g["sk_value"].item()) <= 1e-3 * abs(g["sk_value"].item()), (val.item(), g["sk_value"].item())


def test_emd_dynamic_golden(golden):
    from muscle_b200.loss_multilabel import EMD
    from muscle_b200.torchutils import get_dynamic_crops
    g = golden("emd")
    raw1 = g["raw1"].cuda().requires_grad_()
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(g["raw2"].cuda() * 4, 1), dim=1)
    np.random.seed(g["np_seed"])
    c1, c2, _ = get_dynamic_crops(vw1, g["c1"], vw2.detach(), g["c2"])
    loss = EMD()(c1, c2, mode="dynamic")
    loss.backward()
    assert abs(loss.item() - g["loss"].item()) <= 1e-3 * abs(g["loss"].item()), (loss.item(), g["loss"].item())
    assert _cos(raw1.grad.cpu(), g["g1"]) >= 0.999
    torch.testing.assert_close(raw1.grad.cpu(), g["g1"], rtol=2e-2, atol=2e-3 * float(g["g1"].abs().max()))


def test_emd_generic_lists_and_full_size():
    """Plain python lists of tensors (not produced by our feeder) take the concatenating path; full
    224x224 CAM maps (train_mcl.py sizes) against the oracle."""
    from muscle_b200.loss_multilabel import EMD
    from muscle_b200.torchutils import get_dynamic_crops
    gen = torch.Generator().manual_seed(21)
    b = 2
    raw1 = F.avg_pool2d(torch.randn(b, 21, 224, 224, generator=gen), 9, 1, 4)
    raw2 = F.avg_pool2d(torch.randn(b, 21, 224, 224, generator=gen), 9, 1, 4)
    coord1 = torch.tensor([[0, 30, 180, 150], [40, 0, 120, 200]])
    coord2 = torch.tensor([[44, 0, 180, 150], [0, 24, 120, 200]])
    rr = raw1.clone().requires_grad_()
    v1 = F.normalize(O.cam_softmaxnorm(rr), dim=1)
    v2 = F.normalize(O.cam_softmaxnorm(raw2), dim=1)
    np.random.seed(7)
    r1, r2, _ = O.get_dynamic_crops(v1, coord1, v2.detach(), coord2)
    ref = O.emd_dynamic(r1, r2)
    gref, = torch.autograd.grad(ref, rr)
    rg = raw1.cuda().requires_grad_()
    w1 = F.normalize(O.cam_softmaxnorm(rg), dim=1)
    w2 = F.normalize(O.cam_softmaxnorm(raw2.cuda()), dim=1)
    np.random.seed(7)
    c1, c2, _ = get_dynamic_crops(w1, coord1, w2.detach(), coord2)
    loss = EMD()(c1, c2, mode="dynamic")
    loss.backward()
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item()), (loss.item(), ref.item())
    assert _cos(rg.grad.cpu(), gref) >= 0.999
    # generic path: hand plain lists
    plain1 = [[t.detach().clone().requires_grad_() for t in img] for img in c1]
    plain2 = [[t.detach().clone() for t in img] for img in c2]
    l2 = EMD()(plain1, plain2, mode="dynamic")
    l2.backward()
    assert abs(l2.item() - loss.item()) <= 1e-5 * abs(loss.item())
    assert sum(float(t.grad.abs().sum()) > 0 for img in plain1 for t in img) == len(plain1)   # one crop per image


def _static_views(g, device):
    raw1 = g["raw1"].to(device).requires_grad_()
    vw1 = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    vw2 = F.normalize(torch.softmax(g["raw2"].to(device) * 4, 1), dim=1)
    return raw1, vw1, vw2


def test_emd_static_golden(golden):
    """EMD()(…) in its default mode on get_static_crops windows, against the reference-minted fixture:
    every pair distance, the two-smallest sum and the gradient reaching the raw CAMs."""
    from muscle_b200.loss_multilabel import EMD
    from muscle_b200.torchutils import get_static_crops
    g = golden("emd_static")
    raw1, vw1, vw2 = _static_views(g, "cuda")
    c1, c2, idx = get_static_crops(vw1, g["c1"], vw2.detach(), g["c2"])
    assert idx == list(g["batch_indices"])
    assert [tuple(t.shape) for t in c1] == g["crop_shapes1"]
    assert [tuple(t.shape) for t in c2] == g["crop_shapes2"]
    loss, dists = EMD().static_matching(c1, c2, return_dists=True)
    torch.testing.assert_close(dists.cpu(), g["dists"], rtol=1e-4, atol=1e-6)
    loss.backward()
    assert abs(loss.item() - g["loss"].item()) <= 1e-3 * abs(g["loss"].item()), (loss.item(), g["loss"].item())
    assert _cos(raw1.grad.cpu(), g["g1"]) >= 0.999
    torch.testing.assert_close(raw1.grad.cpu(), g["g1"], rtol=1e-3, atol=1e-4 * float(g["g1"].abs().max()))
    again = EMD()(c1, c2)                       # default mode, as the reference's signature
    assert again.item() == loss.item()          # deterministic


def test_emd_static_matches_oracle_ragged_and_batched():
    """Batched [n,C,h,w] regions of ragged sizes incl. tall, wide, 1-pixel-wide resize targets and
    a tie (duplicated pair): distances vs the oracle, selection = the oracle's stable order."""
    from muscle_b200.loss_multilabel import EMD
    gen = torch.Generator().manual_seed(5)
    shapes = [((3, 16, 20, 9), (3, 16, 7, 7)), ((3, 16, 5, 33), (3, 16, 40, 6)), ((3, 16, 64, 64), (3, 16, 31, 17)),
              ((3, 16, 8, 8), (3, 16, 3, 2)), ((3, 16, 50, 4), (3, 16, 4, 50))]
    a = [F.normalize(torch.rand(s1, generator=gen), dim=1) for s1, _ in shapes]
    b = [F.normalize(torch.rand(s2, generator=gen), dim=1) for _, s2 in shapes]
    a.append(a[1].clone())
    b.append(b[1].clone())                      # equal distances: list.sort() keeps the first
    ra = [t.clone().requires_grad_() for t in a]
    ref, ref_d = O.emd_static(ra, b)
    ref.backward()
    ga = [t.cuda().requires_grad_() for t in a]
    loss, dists = EMD().static_matching(ga, [t.cuda() for t in b], return_dists=True)
    loss.backward()
    torch.testing.assert_close(dists.cpu(), ref_d, rtol=1e-4, atol=1e-6)
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item())
    for ours, theirs in zip(ga, ra):
        assert (ours.grad is None or float(ours.grad.abs().sum()) == 0) == (theirs.grad is None or float(theirs.grad.abs().sum()) == 0)
        if theirs.grad is not None and float(theirs.grad.abs().sum()) > 0:
            torch.testing.assert_close(ours.grad.cpu(), theirs.grad, rtol=1e-3, atol=1e-7)


def test_emd_static_errors_are_loud():
    from muscle_b200.loss_multilabel import EMD
    x = torch.rand(1, 21, 16, 16, device="cuda")
    with pytest.raises(ValueError):
        EMD()([x], [x])                         # fewer than two pairs: the reference raises IndexError
    with pytest.raises(ValueError):
        EMD()([x, x], [x])
    with pytest.raises((ValueError, RuntimeError)):
        EMD()([x.cpu(), x.cpu()], [x.cpu(), x.cpu()])
    with pytest.raises(ValueError):
        EMD()([torch.rand(1, 21, 200, 3, device="cuda")] * 2, [x, x])      # resizes to 0 x 7


def test_static_crops_are_exact_windows_and_generic_lists_agree(golden):
    """get_static_crops windows are bit-exact copies of the python slices the reference takes; plain
    lists of sliced views (not made by our feeder) take the concatenating path and give the same
    loss and the same gradient on the raw maps."""
    from muscle_b200.loss_multilabel import EMD
    from muscle_b200.torchutils import get_static_crops
    g = golden("emd_static")
    raw1, vw1, vw2 = _static_views(g, "cuda")
    c1, c2, _ = get_static_crops(vw1, g["c1"], vw2.detach(), g["c2"])
    r1, r2, _ = O.get_static_crops(vw1.detach(), g["c1"], vw2.detach(), g["c2"])      # slicing on the GPU tensors
    for ours, ref in ((c1, r1), (c2, r2)):
        assert len(ours) == len(ref)
        for t, u in zip(ours, ref):
            assert torch.equal(t.detach(), u)
    loss = EMD()(c1, c2)
    loss.backward()
    g_packed = raw1.grad.clone()
    raw1.grad = None
    vw1b = F.normalize(torch.softmax(raw1 * 4, 1), dim=1)
    p1, p2, _ = O.get_static_crops(vw1b, g["c1"], vw2.detach(), g["c2"])
    l2 = EMD()(list(p1), list(p2))
    l2.backward()
    assert abs(l2.item() - loss.item()) <= 1e-6 * abs(loss.item())
    torch.testing.assert_close(raw1.grad, g_packed, rtol=1e-4, atol=1e-7 * float(g_packed.abs().max()) + 1e-12)
    shuffled = type(c1)(reversed(c1))       # a reordered feeder list must not be read through stale offsets
    shuffled.flat, shuffled.offsets = c1.flat, c1.offsets
    l3 = EMD()(shuffled, type(c2)(reversed(c2)))
    assert abs(l3.item() - loss.item()) <= 1e-5 * abs(loss.item())     # same pairs in another order: same two smallest
