"""Makes the UNMODIFIED reference (laitalaj/cvpce at /root/reference) importable under the installed torch/torchvision.

Authoring-container only: /root/reference does not exist on the GPU box, so nothing that runs there imports this.
Used by make_golden.py (fixtures) and tests/test_reference_pin.py (skipped when the reference tree is absent).

The reference needs matplotlib / squarify / skimage / ray (plotting, hyper-parameter search: not on the hot path and
not installed) and `torchvision.models.utils` (removed upstream).  They are replaced by inert stand-ins BEFORE the
reference is imported; torch itself must be imported first or its own `inspect` walks trip over the stand-ins.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("CVPCE_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "cvpce"))


class _Inert:
    """Absorbs any attribute access / call made at import time by plotting or tuning code."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Inert()

    def __call__(self, *a, **k):
        return _Inert()

    def __iter__(self):
        return iter(())


def _install_stub(name):
    mod = types.ModuleType(name)

    def _getattr(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Inert()

    mod.__getattr__ = _getattr
    sys.modules[name] = mod
    return mod


_loaded = None


def load():
    """Returns the imported reference package `cvpce` (with models.classification, production, datautils loaded)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    import torch  # noqa: F401  (must precede the stubs)
    import torchvision

    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.collections", "matplotlib.offsetbox", "matplotlib.patches",
                 "matplotlib.patheffects", "squarify", "skimage", "skimage.segmentation", "skimage.filters", "ray", "ray.tune"):
        if name not in sys.modules:
            _install_stub(name)
    if "torchvision.models.utils" not in sys.modules:
        from torch.hub import load_state_dict_from_url

        tvu = types.ModuleType("torchvision.models.utils")
        tvu.load_state_dict_from_url = load_state_dict_from_url
        sys.modules["torchvision.models.utils"] = tvu
        torchvision.models.utils = tvu
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import cvpce
        import cvpce.datautils
        import cvpce.models.classification
        import cvpce.production
        import cvpce.utils
    _loaded = cvpce
    return cvpce
