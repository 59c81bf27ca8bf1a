"""`muscle_b200.compat.install` — the three-line alias that makes the reference trainers use the CUDA path
(INTEGRATION.md §3): every name on the path is rebound inside a `src`-shaped package (CPU test on a stand-in
package), and the rebound `MuSCLe.PCM` method and `edge.FieldLoss` compute what the reference's do (GPU test
against the reference-minted fixtures)."""
import random
import sys
import types

import pytest
import torch
import torch.nn.functional as F


def _stand_in_package(name):
    """A package with the reference's module / attribute names (src/__init__.py:1-6 re-exports with *)."""
    pkg = types.ModuleType(name)
    pkg.__path__ = []
    sys.modules[name] = pkg
    mods = {}
    for sub, attrs in (("loss_multilabel", ["info_nce", "image_level_contrast", "PixPro", "EMD", "FocalLoss", "Log_Sum_Exp_Pairwise_Loss"]),
                       ("torchutils", ["get_dynamic_crops", "get_static_crops"]), ("edge", ["FieldLoss"]), ("MuSCLe", [])):
        m = types.ModuleType(name + "." + sub)
        for a in attrs:
            setattr(m, a, ("reference", a))
        sys.modules[name + "." + sub] = m
        setattr(pkg, sub, m)
        mods[sub] = m

    class MuSCLe(torch.nn.Module):           # the reference model keeps PCM as a method using self.fuse
        def __init__(self, cin=208):
            super().__init__()
            self.fuse = torch.nn.Conv2d(cin, 128, 1, bias=True)

        def PCM(self, cam, f):
            raise AssertionError("the reference method should have been rebound")

    mods["MuSCLe"].MuSCLe = MuSCLe
    for a in ("image_level_contrast", "PixPro", "EMD", "FieldLoss"):   # `from src import *`
        setattr(pkg, a, ("reference", a))
    return pkg, mods


def _drop(name):
    for k in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
        del sys.modules[k]


def test_install_rebinds_every_name_on_the_path():
    from muscle_b200 import compat, edge, loss_multilabel, torchutils
    pkg, mods = _stand_in_package("fake_src_cpu")
    try:
        done = compat.install("fake_src_cpu")
        assert set(done) == {"loss_multilabel.info_nce", "loss_multilabel.image_level_contrast", "loss_multilabel.PixPro",
                             "loss_multilabel.EMD", "loss_multilabel.FocalLoss", "loss_multilabel.Log_Sum_Exp_Pairwise_Loss",
                             "torchutils.get_dynamic_crops", "torchutils.get_static_crops",
                             "edge.FieldLoss", "MuSCLe.MuSCLe.PCM"}
        assert mods["loss_multilabel"].info_nce is loss_multilabel.info_nce
        assert mods["loss_multilabel"].image_level_contrast is loss_multilabel.image_level_contrast
        assert mods["loss_multilabel"].PixPro is loss_multilabel.PixPro
        assert mods["loss_multilabel"].EMD is loss_multilabel.EMD
        assert mods["torchutils"].get_dynamic_crops is torchutils.get_dynamic_crops
        assert mods["torchutils"].get_static_crops is torchutils.get_static_crops
        assert mods["edge"].FieldLoss is edge.FieldLoss
        assert pkg.FieldLoss is edge.FieldLoss and pkg.EMD is loss_multilabel.EMD      # the * re-exports too
        assert pkg.image_level_contrast is loss_multilabel.image_level_contrast and pkg.PixPro is loss_multilabel.PixPro
        model = mods["MuSCLe"].MuSCLe()
        with pytest.raises(RuntimeError):        # rebound: reaches the CUDA path, which refuses CPU tensors loudly
            model.PCM(torch.rand(1, 21, 4, 4), torch.rand(1, 208, 4, 4))
    finally:
        _drop("fake_src_cpu")


@pytest.mark.gpu
def test_rebound_pcm_and_fieldloss_match_the_reference(golden):
    from muscle_b200 import compat
    pkg, mods = _stand_in_package("fake_src_gpu")
    try:
        compat.install("fake_src_gpu")
        g = golden("heads")
        cin = g["fuse_w"].shape[1]
        model = mods["MuSCLe"].MuSCLe(cin).cuda()
        with torch.no_grad():
            model.fuse.weight.copy_(g["fuse_w"])
            model.fuse.bias.copy_(g["fuse_b"])
        cam = g["pcm_cam"].cuda().requires_grad_()
        out = model.PCM(cam, g["pcm_f"].cuda())                 # the call src/MuSCLe.py:254 makes
        (out * g["pcm_go"].cuda()).sum().backward()
        ref = g["pcm_out"]
        assert float((out.detach().cpu() - ref).abs().max() / ref.abs().max()) < 2e-2
        cos = lambda a, b: F.cosine_similarity(a.flatten().double(), b.flatten().double(), dim=0).item()
        assert cos(cam.grad.cpu(), g["pcm_gcam"]) >= 0.999
        assert cos(model.fuse.weight.grad.cpu(), g["pcm_gw"]) >= 0.999
        assert cos(model.fuse.bias.grad.cpu(), g["pcm_gb"]) >= 0.999

        f = golden("fieldloss")
        crit = mods["edge"].FieldLoss(sobel_size=5, beta=f["beta"], k=f["k"])     # train_muscle.py:163
        dense = f["dense"].cuda().requires_grad_()
        random.seed(f["py_seed"])
        loss, mag = crit(f["seg"].cuda(), dense, f["mask"].cuda(), f["label_bg"].cuda(), f["step"])   # :196
        assert torch.is_tensor(loss)
        loss.backward()
        assert abs(loss.item() - f["loss"].item()) <= 1e-3 * abs(f["loss"].item())
        assert cos(dense.grad.cpu(), f["gd"]) >= 0.999
    finally:
        _drop("fake_src_gpu")
