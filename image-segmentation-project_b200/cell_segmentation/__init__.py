"""`cell_segmentation` package shim (reference main.py:5,9,13,17 imports
`cell_segmentation.<experiment module>`; the reference repository itself is that package's
directory, checked out under that name -- SURVEY F4).

`cell_segmentation.X` resolves to the top-level module `X` found on sys.path: the B200 modules of
this directory (`advanced_models`, `losses`, `utils`, `train`, `dataset`, `cross_validation`) and
the reference's experiment scripts (`forward_selection_integration`, `loss_function_cross_validation`,
`model_configuration_experiment`, `augmentation_amount_experiment`) from wherever the user keeps
them.  With this directory first on PYTHONPATH, `python main.py --experiment ...` runs unchanged.
"""
import importlib
import importlib.abc
import importlib.util
import sys


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    prefix = __name__ + "."

    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(self.prefix) or "." in fullname[len(self.prefix):]:
            return None
        return importlib.util.spec_from_loader(fullname, self)

    def create_module(self, spec):
        return importlib.import_module(spec.name[len(self.prefix):])

    def exec_module(self, module):
        pass


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    try:
        return importlib.import_module(__name__ + "." + name)
    except ImportError as e:
        raise AttributeError(f"cell_segmentation has no module {name!r}: {e}") from None
