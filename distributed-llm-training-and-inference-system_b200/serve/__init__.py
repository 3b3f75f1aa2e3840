"""Serving layer (reference llmctl/serve/__init__.py:3-5 exports InferenceServer, create_inference_server)."""


def __getattr__(name):
    if name in ("InferenceServer", "create_inference_server", "GenerationRequest", "GenerationResponse"):
        from . import server
        return getattr(server, name)
    if name in ("InferenceEngine", "load_model_for_inference"):
        from . import engine
        return getattr(engine, name)
    if name in ("KVBlockManager", "BatchScheduler"):
        from . import native_host
        return getattr(native_host, name)
    raise AttributeError(name)
