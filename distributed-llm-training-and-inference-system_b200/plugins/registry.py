"""Plugin registry for the four groups the reference declares but never ships: ``kernels``, ``quantizers``,
``exporters``, ``schedulers`` (entry-point group ``llmctl.plugins`` in reference pyproject.toml:77-81 points
at modules that do not exist; the contract — a module-level ``register()`` returning ``{name: class}``,
schedulers exposing ``.schedule(requests)`` — is sketched in reference docs/USER_GUIDE.md:260-281).

Plugins are discovered from (1) installed entry points of group ``llmctl.plugins`` whose name is one of the
four groups, (2) modules named in ``LLMCTL_PLUGINS=group:module[,group:module...]``, (3) direct calls to
``registry.register(group, name, obj)``.  Built-ins register the native kernels and the C++ scheduler.
"""
from __future__ import annotations

import importlib
import os
from typing import Any, Callable, Dict, Optional

GROUPS = ("kernels", "quantizers", "exporters", "schedulers")


class PluginRegistry:
    def __init__(self):
        self._items: Dict[str, Dict[str, Any]] = {g: {} for g in GROUPS}
        self._loaded = False

    def register(self, group: str, name: str, obj: Any, overwrite: bool = False) -> None:
        if group not in self._items:
            raise KeyError(f"unknown plugin group {group!r}; groups: {GROUPS}")
        if name in self._items[group] and not overwrite:
            raise ValueError(f"{group} plugin {name!r} already registered")
        self._items[group][name] = obj

    def register_module(self, group: str, module_name: str) -> int:
        mod = importlib.import_module(module_name)
        reg: Optional[Callable[[], Dict[str, Any]]] = getattr(mod, "register", None)
        if reg is None:
            raise AttributeError(f"plugin module {module_name} has no register()")
        items = reg()
        for name, obj in items.items():
            self.register(group, name, obj, overwrite=True)
        return len(items)

    def _discover(self) -> None:
        if self._loaded:
            return
        self._loaded = True
        _register_builtins(self)
        try:
            from importlib.metadata import entry_points
            for ep in entry_points(group="llmctl.plugins"):
                if ep.name in GROUPS:
                    try:
                        items = ep.load()()
                        for name, obj in items.items():
                            self.register(ep.name, name, obj, overwrite=True)
                    except Exception:  # noqa: BLE001 - a broken third-party plugin must not break the CLI
                        pass
        except Exception:  # noqa: BLE001
            pass
        for spec in filter(None, os.environ.get("LLMCTL_PLUGINS", "").split(",")):
            group, _, module = spec.partition(":")
            try:
                self.register_module(group.strip(), module.strip())
            except Exception:  # noqa: BLE001
                pass

    def get(self, group: str, name: str) -> Optional[Any]:
        self._discover()
        return self._items.get(group, {}).get(name)

    def list(self, group: Optional[str] = None) -> Dict[str, Any]:
        self._discover()
        if group:
            return dict(self._items[group])
        return {g: sorted(v) for g, v in self._items.items()}


def _register_builtins(reg: PluginRegistry) -> None:
    from ..ops import functional as ops
    for name in ("rmsnorm", "layernorm", "rope_packed", "swiglu", "linear", "cross_entropy", "attention", "gemm_raw"):
        reg.register("kernels", name, getattr(ops, name), overwrite=True)

    class NativeScheduler:
        """Adapter with the user-guide contract: ``schedule(requests) -> batch`` (requests need
        ``.request_id``, ``.prompt_len``, ``.max_tokens``)."""

        def __init__(self, max_batch_size: int = 8, max_batch_tokens: int = 8192, policy: str = "dynamic"):
            from ..serve.native_host import BatchScheduler
            self.core = BatchScheduler(max_batch_size, max_batch_tokens, policy)

        def schedule(self, requests):
            by_id = {}
            for r in requests:
                rid = int(getattr(r, "request_id", id(r)) if not isinstance(r, dict) else r["request_id"])
                pl = getattr(r, "prompt_len", None) if not isinstance(r, dict) else r["prompt_len"]
                mt = getattr(r, "max_tokens", None) if not isinstance(r, dict) else r["max_tokens"]
                if self.core.add_request(rid, int(pl), int(mt)):
                    by_id[rid] = r
            return [by_id[i] for i, _, _ in self.core.step() if i in by_id]

    for pol in ("dynamic", "simple", "vllm-like"):
        reg.register("schedulers", pol, (lambda p: (lambda **kw: NativeScheduler(policy=p, **kw)))(pol), overwrite=True)

    class SafetensorsExporter:
        def export(self, ckpt, out, quant=None, **kw):
            from ..io.export import export_checkpoint
            return export_checkpoint(ckpt, out, "safetensors", quant, **kw)

    reg.register("exporters", "safetensors", SafetensorsExporter, overwrite=True)

    from ..io.quantize import BUILTIN as _QUANT
    for name, cls in _QUANT.items():
        reg.register("quantizers", name, cls, overwrite=True)


_REGISTRY: Optional[PluginRegistry] = None


def get_registry() -> PluginRegistry:
    global _REGISTRY
    if _REGISTRY is None:
        _REGISTRY = PluginRegistry()
    return _REGISTRY
