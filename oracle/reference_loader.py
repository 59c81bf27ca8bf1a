"""Import the UNMODIFIED reference from /root/reference (authoring container only).

TEST INFRASTRUCTURE ONLY.  The reference's ``import src`` pulls in packages that are
not installed here and that no hot-path function uses: they are
replaced by empty stub modules before the import.  Nothing is copied; the modules
are executed from where they lie.  The GPU box has no /root/reference, so only
``oracle/make_golden.py`` (run here) uses this loader; tests use the committed
fixtures instead.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("MUSCLE_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src"))


def load():
    """Returns the reference ``src`` package (loss_multilabel, edge, torchutils, MuSCLe…)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    stubs = {
        "qpth": {},
        "qpth.qp": {"QPFunction": object},      # only used by the uncalled emd_inference_qpth
        "imageio": {},
        "skimage": {},
        "skimage.transform": {"resize": None},
    }
    for name, attrs in stubs.items():
        if name not in sys.modules:
            mod = types.ModuleType(name)
            for k, v in attrs.items():
                setattr(mod, k, v)
            sys.modules[name] = mod
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import src  # noqa: E402  (the reference package)
        import src.efficientnet_pytorch.model as effmodel
        import src.torchutils, src.edge, src.loss_multilabel, src.MuSCLe  # noqa: F401
        # constructing MuSCLe(...) would download weights; make that a no-op
        effmodel.load_pretrained_weights = lambda *a, **k: None
    return src
