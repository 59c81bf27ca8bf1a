"""Aliases the CUDA path into an imported reference tree (`src`), so train_mcl.py / train_muscle.py
run unmodified.  Only the names on the contrastive-loss path are rebound."""
from __future__ import annotations

import importlib


def install(src_package: str = "src"):
    """Rebinds, inside the already importable reference package `src_package`:
    loss_multilabel.{info_nce,image_level_contrast,PixPro,EMD,FocalLoss,Log_Sum_Exp_Pairwise_Loss}, torchutils.{get_dynamic_crops,get_static_crops},
    edge.FieldLoss, and MuSCLe.PCM.  Returns the list of rebound names."""
    from . import edge, heads, loss_multilabel, torchutils

    done = []

    def rebind(mod_name, attr, value):
        mod = importlib.import_module(src_package + "." + mod_name)
        setattr(mod, attr, value)
        done.append(mod_name + "." + attr)

    rebind("loss_multilabel", "info_nce", loss_multilabel.info_nce)
    rebind("loss_multilabel", "image_level_contrast", loss_multilabel.image_level_contrast)
    rebind("loss_multilabel", "PixPro", loss_multilabel.PixPro)
    rebind("loss_multilabel", "EMD", loss_multilabel.EMD)
    rebind("loss_multilabel", "FocalLoss", loss_multilabel.FocalLoss)
    rebind("loss_multilabel", "Log_Sum_Exp_Pairwise_Loss", loss_multilabel.Log_Sum_Exp_Pairwise_Loss)
    rebind("torchutils", "get_dynamic_crops", torchutils.get_dynamic_crops)
    rebind("torchutils", "get_static_crops", torchutils.get_static_crops)
    rebind("edge", "FieldLoss", edge.FieldLoss)
    top = importlib.import_module(src_package)
    for name, value in (("image_level_contrast", loss_multilabel.image_level_contrast), ("PixPro", loss_multilabel.PixPro),
                        ("EMD", loss_multilabel.EMD), ("FieldLoss", edge.FieldLoss), ("FocalLoss", loss_multilabel.FocalLoss),
                        ("Log_Sum_Exp_Pairwise_Loss", loss_multilabel.Log_Sum_Exp_Pairwise_Loss)):
        if hasattr(top, name):           # `from src import *` re-exports (src/__init__.py:1-6)
            setattr(top, name, value)
    muscle_mod = importlib.import_module(src_package + ".MuSCLe")

    def PCM(self, cam, f):               # src/MuSCLe.py:213-223, a method using self.fuse
        return heads.pcm(cam, f, self.fuse.weight, self.fuse.bias)

    muscle_mod.MuSCLe.PCM = PCM
    done.append("MuSCLe.MuSCLe.PCM")
    return done
