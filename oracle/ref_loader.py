"""Import the reference's OWN modules from oracle/_ref (staged by oracle/make_ref.py) -- test and
CPU-baseline infrastructure only.

`reference_modules()` returns the reference's `advanced_models`, `losses`, `utils`, `train` module
objects, imported under their flat names with oracle/_ref FIRST on sys.path, exactly as the
reference's scripts import each other (train.py:13-14).  matplotlib / seaborn / albumentations are
absent on both machines and sit only on the plotting / augmentation side: they are stubbed at
import time (classes that record their constructor arguments and raise when called).
Must be called in a process that has NOT imported the B200 modules of the same names.
"""
import importlib
import os
import sys
import types

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_FLAT = ("dataset", "utils", "advanced_models", "losses", "train", "cross_validation")


def available():
    return os.path.exists(os.path.join(REF_DIR, "train.py"))


def _stub(name):
    m = types.ModuleType(name)

    class _Recorded:
        def __init__(self, *a, **k):
            self.args, self.kwargs = a, k

        def __call__(self, *a, **k):
            raise RuntimeError(f"{name} is not installed (import-time stub)")

    def _getattr(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Recorded

    m.__getattr__ = _getattr
    sys.modules[name] = m


def reference_modules():
    if not available():
        raise FileNotFoundError(f"{REF_DIR} is empty: run `python oracle/make_ref.py` in the build container")
    clash = [n for n in _FLAT if n in sys.modules and
             not os.path.abspath(getattr(sys.modules[n], "__file__", "") or "").startswith(REF_DIR)]
    if clash:
        raise RuntimeError(f"modules {clash} are already imported from elsewhere; use a fresh process")
    for n in ("albumentations", "matplotlib", "matplotlib.pyplot", "seaborn"):
        if n not in sys.modules:
            try:
                if importlib.util.find_spec(n.split(".")[0]) is not None and n.split(".")[0] != "albumentations":
                    continue
            except (ImportError, ValueError):
                pass
            _stub(n)
    if REF_DIR in sys.path:
        sys.path.remove(REF_DIR)
    sys.path.insert(0, REF_DIR)
    mods = {n: importlib.import_module(n) for n in ("advanced_models", "losses", "utils", "train")}
    return types.SimpleNamespace(**mods)
