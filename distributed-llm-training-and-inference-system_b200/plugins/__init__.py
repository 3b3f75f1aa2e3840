"""Plugins layer — exports of reference llmctl/plugins/__init__.py:3-9 plus the registry."""
from .autotuning import AutoTuner, Tunable, TuningConfig, TuningResult, create_auto_tuner  # noqa: F401
from .registry import GROUPS, PluginRegistry, get_registry  # noqa: F401
