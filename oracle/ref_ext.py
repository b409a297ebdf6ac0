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


# ---- the reference's own Python wrappers on top of a given `_backend` -----------------------------------------------------
_WRAPPER_FILES = {"gridencoder": "grid.py", "raymarching": "raymarching.py", "ffmlp": "ffmlp.py", "shencoder": "sphere_harmonics.py"}
_BACKEND_NAMES = {"gridencoder": "_gridencoder", "raymarching": "_raymarching", "ffmlp": "_ffmlp", "shencoder": "_shencoder"}
# imports of the reference modules that have nothing to do with the path (and are not installed here): stubbed for the import
_STUB_MODULES = ("MinkowskiEngine", "matplotlib", "matplotlib.pyplot", "turtle")
_STUB_RELATIVE = ("resfieldnet", "minkvae", "minkfieldunet", "style_unet", "stylegan1_3d", "stylegan2_3d")


def load_wrapper(pkg, backend):
    """Import the reference's L1 module of `pkg` (staged by build_ref_ext.stage_wrappers into oracle/_ref/py/) with
    `backend` standing in for its pybind11 extension (`import _gridencoder as _backend`, grid.py:22-25 and alike).
    Returns the module, or None when the wrappers were not staged."""
    import sys
    import types
    key = ("wrapper", pkg, id(backend))
    if key in _cache:
        return _cache[key]
    path = os.path.join(_HERE, "_ref", "py", pkg, _WRAPPER_FILES[pkg])
    if not os.path.exists(path):
        return None

    class _Anything(types.ModuleType):  # `from .resfieldnet import ResNet50` and friends: any name resolves to a dummy class
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return type(name, (), {})

    pkg_name = f"ngp_refpy_{pkg}"
    parent = types.ModuleType(pkg_name)
    parent.__path__ = [os.path.dirname(path)]
    injected = {pkg_name: parent, _BACKEND_NAMES[pkg]: backend}
    for r in _STUB_RELATIVE:
        injected[f"{pkg_name}.{r}"] = _Anything(f"{pkg_name}.{r}")
    for m in _STUB_MODULES:
        try:
            __import__(m)
        except Exception:  # noqa: BLE001 -- not installed (matplotlib) or unusable (turtle needs tkinter)
            injected[m] = _Anything(m)
    if "matplotlib" in injected:
        injected["matplotlib"].use = lambda *a, **k: None
    saved = {k: sys.modules.get(k) for k in injected}
    sys.modules.update(injected)
    try:
        spec = importlib.util.spec_from_file_location(f"{pkg_name}.{_WRAPPER_FILES[pkg][:-3]}", path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if k == pkg_name:
                continue  # keep the synthetic parent: pickling / repr of the classes refer to it
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    assert mod._backend is backend, "the reference wrapper did not pick up the given backend"
    _cache[key] = mod
    return mod
