"""Import shim: registers the package directory `4paradigm-k8s-device-plugin_b200/`
(not a valid Python identifier) under the importable name ``vgpu_b200``."""
import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG_DIR = os.path.join(_HERE, "4paradigm-k8s-device-plugin_b200")


def _load():
    if "vgpu_b200" in sys.modules and getattr(sys.modules["vgpu_b200"], "__path__", None):
        return sys.modules["vgpu_b200"]
    spec = importlib.util.spec_from_file_location(
        "vgpu_b200", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR]
    )
    mod = importlib.util.module_from_spec(spec)
    sys.modules["vgpu_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


_load()
