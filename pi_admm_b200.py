"""Import alias: the package directory is named after the reference repository
(``distributed-local-planner-pi-admm_b200``), which is not a valid Python identifier.
``import pi_admm_b200`` and ``import pi_admm_b200.<submodule>`` resolve to the very same module
objects as the hyphenated name (one copy of every module, hence one set of ctypes types)."""
import importlib
import importlib.abc
import importlib.util
import os
import sys

_REAL = "distributed-local-planner-pi-admm_b200"
_ALIAS = "pi_admm_b200"
_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real):
        self.real = real

    def create_module(self, spec):
        return importlib.import_module(self.real)

    def exec_module(self, module):
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname == _ALIAS or fullname.startswith(_ALIAS + "."):
            real = _REAL + fullname[len(_ALIAS):]
            return importlib.util.spec_from_loader(fullname, _AliasLoader(real))
        return None


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())
_pkg = importlib.import_module(_REAL)
sys.modules[__name__] = _pkg
