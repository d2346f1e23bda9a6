"""Put this directory FIRST on PYTHONPATH and set B200ATTN_DROPIN=1 to apply the drop-in in every
process, including the tensor-parallel workers the reference spawns (llm_engine.py:103-106 uses the
"spawn" start method, so a patch made in the parent is not inherited).  An import hook calls
`b200attn.install.install()` right after `nanovllm.engine.model_runner` is first imported.

B200ATTN_DROPIN=attn limits the patch to attention + KV scatter (no pointwise, no ModelRunner)."""
import importlib.abc
import importlib.machinery
import os
import sys

_MODE = os.environ.get("B200ATTN_DROPIN", "")
_TRIGGER = "nanovllm.engine.model_runner"


class _Loader(importlib.abc.Loader):
    def __init__(self, inner):
        self.inner = inner

    def create_module(self, spec):
        return self.inner.create_module(spec)

    def exec_module(self, module):
        self.inner.exec_module(module)
        here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        if here not in sys.path:
            sys.path.insert(0, here)
        from b200attn.install import install
        full = _MODE != "attn"
        install(pointwise=full, runner=full)


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
        if name != _TRIGGER:
            return None
        sys.meta_path.remove(self)
        try:
            spec = importlib.machinery.PathFinder.find_spec(name, path)
        finally:
            pass
        if spec is not None and spec.loader is not None:
            spec.loader = _Loader(spec.loader)
        return spec


if _MODE and _MODE != "0":
    sys.meta_path.insert(0, _Finder())
