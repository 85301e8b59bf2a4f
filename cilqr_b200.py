"""Import alias: ``import cilqr_b200`` is the package directory
``uncertainty-aware-cilqr-for-trajectory-optimization_b200/`` (whose name is not a Python identifier).
Sub-modules are aliased too, so ``cilqr_b200.solver`` and the long name are the same module objects."""
import importlib
import sys

_REAL = "uncertainty-aware-cilqr-for-trajectory-optimization_b200"
_real = importlib.import_module(_REAL)
for _name, _mod in list(sys.modules.items()):
    if _name.startswith(_REAL + "."):
        sys.modules[__name__ + _name[len(_REAL):]] = _mod
sys.modules[__name__] = _real
