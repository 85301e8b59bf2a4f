"""Import alias: ``import cilqr_b200`` == the package directory
``uncertainty-aware-cilqr-for-trajectory-optimization_b200/`` (whose name is not a Python identifier)."""
import importlib
import sys

sys.modules[__name__] = importlib.import_module("uncertainty-aware-cilqr-for-trajectory-optimization_b200")
