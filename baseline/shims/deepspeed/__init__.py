"""Import shim: the reference imports ``deepspeed`` (reference llmctl/runtime/engine.py:25) and never uses it.
The real package is not in this image (no network).  Empty on purpose."""
__version__ = "0.0.0"
