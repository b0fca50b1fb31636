"""Imports the package directory `eradiate-kernel_b200/` (not a valid Python identifier)
under the module name `eradiate_kernel_b200`."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
NAME = "eradiate_kernel_b200"


def load():
    if NAME in sys.modules:
        return sys.modules[NAME]
    pkg = os.path.join(ROOT, "eradiate-kernel_b200")
    spec = importlib.util.spec_from_file_location(NAME, os.path.join(pkg, "__init__.py"),
                                                  submodule_search_locations=[pkg])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[NAME] = mod
    spec.loader.exec_module(mod)
    return mod
