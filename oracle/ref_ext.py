"""TEST INFRASTRUCTURE ONLY -- loads the reference's own CUDA extensions prebuilt by
oracle/build_ref_ext.py (oracle/_ref/_<name>.so).  Returns None when a module is absent."""
import importlib.machinery
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_cache = {}


def load(name):
    """name in {'_gridencoder', '_raymarching', '_ffmlp', '_shencoder'}"""
    if name in _cache:
        return _cache[name]
    path = os.path.join(_HERE, "_ref", name + ".so")
    mod = None
    if os.path.exists(path):
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        loader = importlib.machinery.ExtensionFileLoader(name, path)
        spec = importlib.util.spec_from_loader(name, loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
    _cache[name] = mod
    return mod
