"""Imports the package directory `flink-tensorflow_b200/` (not a valid module name) as `flink_tensorflow_b200`."""
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))


def load():
    name = "flink_tensorflow_b200"
    if name in sys.modules:
        return sys.modules[name]
    pkg_dir = os.path.join(_ROOT, "flink-tensorflow_b200")
    spec = importlib.util.spec_from_file_location(name, os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    try:
        spec.loader.exec_module(mod)
    except BaseException:
        sys.modules.pop(name, None)
        raise
    return mod
