"""Import alias: the package source lives in ``distributed-llm-training-and-inference-system_b200/``
(a directory name that is not a valid Python identifier).  Importing ``llmctl_b200`` executes that
directory's ``__init__.py`` with ``__path__`` pointing at it, so ``llmctl_b200.models`` etc. resolve
to the real sub-packages."""
import os as _os

_here = _os.path.dirname(_os.path.abspath(__file__))
_real = _os.path.join(_os.path.dirname(_here), "distributed-llm-training-and-inference-system_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r", encoding="utf-8") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
