"""Importable alias of the package directory ``drone-oa-iree-vit-accelerator_b200/`` (its name is not a Python identifier)."""
import importlib
import sys

_pkg = importlib.import_module("drone-oa-iree-vit-accelerator_b200")
sys.modules[__name__] = _pkg
