"""Entry-point targets for the four plugin groups (pyproject ``llmctl.plugins``): each returns ``{name: obj}``
exactly like a third-party ``register()`` would (reference pyproject.toml:77-81 names modules that do not
exist)."""
from .registry import PluginRegistry, _register_builtins


def _group(g):
    r = PluginRegistry()
    r._loaded = True
    _register_builtins(r)
    return r.list(g)


def register_kernels():
    return _group("kernels")


def register_schedulers():
    return _group("schedulers")


def register_exporters():
    return _group("exporters")


def register_quantizers():
    return _group("quantizers")
