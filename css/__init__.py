"""``css`` -- import alias of ``css_b200`` under the reference's package name.

The reference's scripts import ``css.executor.separator``, ``css.models``, ``css.objectives`` ... (e.g.
scripts/python/separate_libricss.py:14-19, scripts/python/train.py:16-22).  With this package on the path those imports
resolve to the B200-native implementation: ``css.X`` IS the module object ``css_b200.X`` (one copy of every module, so
classes and the loaded shared library are shared between the two names).
"""
import importlib
import importlib.abc
import importlib.util
import sys

import css_b200

_PREFIX = __name__ + "."


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(_PREFIX):
            return None
        real = "css_b200." + fullname[len(_PREFIX):]
        try:
            importlib.import_module(real)
        except ModuleNotFoundError as e:
            if e.name and real.startswith(e.name):
                return None
            raise
        return importlib.util.spec_from_loader(fullname, self)

    def create_module(self, spec):
        mod = sys.modules["css_b200." + spec.name[len(_PREFIX):]]
        self._real[id(mod)] = (mod.__spec__, getattr(mod, "__loader__", None))
        return mod

    def exec_module(self, module):
        # the import system stamped the alias spec on the (shared) module object: put the real one back
        module.__spec__, module.__loader__ = self._real.pop(id(module))

    _real = {}


sys.meta_path.insert(0, _AliasFinder())
__version__ = css_b200.__version__ if hasattr(css_b200, "__version__") else "0"
__path__ = []          # a package with no files of its own: every submodule comes from the finder above


def __getattr__(name):
    try:
        return importlib.import_module(_PREFIX + name)
    except ModuleNotFoundError:
        raise AttributeError(f"module 'css' has no attribute '{name}'") from None
