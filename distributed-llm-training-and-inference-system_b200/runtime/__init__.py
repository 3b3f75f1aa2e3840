"""Runtime layer (reference llmctl/runtime/__init__.py:3-13 re-exports the same names)."""
from .launcher import (BaseLauncher, LaunchConfig, LocalLauncher, MPILauncher, ProcessOrchestrator,  # noqa: F401
                       SlurmLauncher, create_launcher)


def __getattr__(name):
    if name in ("TrainingEngine", "TrainingConfig", "create_training_config"):
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)
