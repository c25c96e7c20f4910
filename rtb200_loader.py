"""Imports the package directory `ray-tracing_b200/` (hyphenated, not a Python identifier) as module `ray_tracing_b200`."""
import importlib.util
import os
import sys

_NAME = "ray_tracing_b200"


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    here = os.path.dirname(os.path.abspath(__file__))
    pkg = os.path.join(here, "ray-tracing_b200")
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(pkg, "__init__.py"), submodule_search_locations=[pkg])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
