"""Import alias for the product package.

The package directory is named after the reference repository
(`research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200/`),
which is not a valid Python identifier.  `import canary_b200` loads that directory
as the package `canary_b200` so that `from canary_b200 import operator_api` works.
"""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(
    os.path.dirname(os.path.abspath(__file__)),
    "research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200")

_spec = importlib.util.spec_from_file_location(
    "canary_b200", os.path.join(_PKG_DIR, "__init__.py"),
    submodule_search_locations=[_PKG_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["canary_b200"] = _mod
_spec.loader.exec_module(_mod)
