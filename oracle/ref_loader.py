"""TEST INFRASTRUCTURE ONLY — imports the UNMODIFIED reference from /root/reference.

Only usable in the build container (the GPU box has no /root/reference).  Used by
tests/golden/make_golden.py to generate the committed golden fixtures and to
validate oracle/amd_oracle.py (the restatement that does travel).

`model/__init__.py` of the reference pulls model_AE/model_MAE (timm, lpips weights),
so `model` is registered as a bare namespace package over /root/reference/model and
the hot-path files are imported directly.
"""
import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("HIVAE_REFERENCE", "/root/reference")
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shim")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "model"))


def load():
    """Returns a namespace with the reference modules: wavelet, frequency_utils,
    attention, rectified_flow, modules, transformer, model_AMD."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    if _SHIM not in sys.path:
        sys.path.insert(0, _SHIM)
    if "model" not in sys.modules or not hasattr(sys.modules["model"], "__hivae_ref__"):
        pkg = types.ModuleType("model")
        pkg.__path__ = [os.path.join(REF_ROOT, "model")]
        pkg.__hivae_ref__ = True
        import importlib.machinery
        pkg.__spec__ = importlib.machinery.ModuleSpec("model", None, is_package=True)
        pkg.__spec__.submodule_search_locations = pkg.__path__
        sys.modules["model"] = pkg
    ns = types.SimpleNamespace()
    for name in ("wavelet", "frequency_utils", "attention", "rectified_flow", "modules", "transformer", "model_AMD"):
        setattr(ns, name, importlib.import_module("model." + name))
    return ns
