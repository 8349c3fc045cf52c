"""Alias: `import b200match` == importlib.import_module("3dreconstruction_b200")."""
import importlib
import sys

sys.modules[__name__] = importlib.import_module("3dreconstruction_b200")
