"""HTTP inference server (FastAPI + uvicorn): ``POST /v1/completions``, ``GET /v1/models``, ``GET /health``,
``GET /metrics`` — the API of reference llmctl/serve/server.py:253-407 (same request / response models,
503 "Server at capacity", CORS ``*``) with the defects fixed: generation runs to EOS or
``max_tokens`` instead of stopping after one token (D1); there is no padding, so no sampling from pad
positions (D2); the engine loop runs on its own thread so a forward never blocks the event loop (D3);
``stream=true`` streams server-sent events instead of being ignored.
"""
from __future__ import annotations

import asyncio
import json
import threading
import time
import uuid
from typing import Dict, List, Optional

from pydantic import BaseModel, Field

from .engine import InferenceEngine, Sequence


class GenerationRequest(BaseModel):
    prompt: str
    max_tokens: int = Field(100, ge=1)
    temperature: float = Field(1.0, ge=0.0)
    top_p: float = Field(1.0, gt=0.0, le=1.0)
    top_k: int = -1
    stream: bool = False


class GenerationResponse(BaseModel):
    id: str
    text: str
    finish_reason: str
    usage: Dict[str, int]


class RequestState:
    """Kept for API parity with reference llmctl/serve/server.py:39-55."""

    def __init__(self, request_id: str, seq: Sequence):
        self.request_id, self.seq = request_id, seq

    @property
    def finished(self) -> bool:
        return self.seq.finished

    @property
    def generated_tokens(self) -> List[int]:
        return self.seq.generated


class EngineLoop(threading.Thread):
    """Drives ``engine.step()`` off the event loop; sleeps on a condition when idle."""

    def __init__(self, engine: InferenceEngine):
        super().__init__(daemon=True, name="llmctl-engine")
        self.engine = engine
        self.wake = threading.Event()
        self.stop_flag = False
        self.error: Optional[BaseException] = None
        self.on_finished = None

    def run(self):
        while not self.stop_flag:
            if not self.engine.has_work():
                self.wake.wait(timeout=0.05)
                self.wake.clear()
                if not self.engine.has_work():
                    continue
            try:
                done = self.engine.step()
                if self.on_finished:
                    for s in done:
                        self.on_finished(s)
            except BaseException as e:  # noqa: BLE001 - fail the in-flight requests, keep serving
                self.error = e
                for s in list(self.engine.seqs.values()):
                    self.engine.abort(s, "error")

    def stop(self):
        self.stop_flag = True
        self.wake.set()


class InferenceServer:
    def __init__(self, model_path: str, host: str = "0.0.0.0", port: int = 8080, max_batch_size: int = 8,
                 max_batch_tokens: int = 8192, max_concurrent_requests: int = 128, scheduler: str = "dynamic",
                 device: Optional[str] = None, engine: Optional[InferenceEngine] = None):
        from fastapi import FastAPI, HTTPException
        from fastapi.middleware.cors import CORSMiddleware
        from fastapi.responses import PlainTextResponse, StreamingResponse

        self.model_path, self.host, self.port = str(model_path), host, port
        self.max_concurrent_requests = max_concurrent_requests
        self.engine = engine or InferenceEngine(model_path, max_batch_size=max_batch_size,
                                                max_batch_tokens=max_batch_tokens, scheduler=scheduler, device=device)
        self.scheduler = self.engine.sched
        self.active_requests: Dict[str, RequestState] = {}
        self.completed = 0
        self.errors = 0
        self.loop_thread = EngineLoop(self.engine)
        self.loop_thread.on_finished = self._record
        if hasattr(self.engine, "on_pending"):            # tensor-parallel serving (serve/lockstep.py)
            self.engine.on_pending = self.loop_thread.wake.set
        try:
            # register the llmctl_inference_* metrics in the default registry: GET /metrics serves them from this
            # process, no second HTTP server is started
            from ..metrics.observability import get_observability_manager
            get_observability_manager().enable_prometheus(port=None)
        except Exception:  # noqa: BLE001 - metrics are optional (prometheus_client missing)
            pass
        self.started = time.time()
        self.app = FastAPI(title="llmctl inference server", version="0.1.0")
        self.app.add_middleware(CORSMiddleware, allow_origins=["*"], allow_credentials=True, allow_methods=["*"],
                                allow_headers=["*"])
        app = self.app

        @app.post("/v1/completions")
        async def create_completion(req: GenerationRequest):
            if len(self.active_requests) >= self.max_concurrent_requests:
                raise HTTPException(status_code=503, detail="Server at capacity")
            if req.stream:
                return StreamingResponse(self._stream(req), media_type="text/event-stream")
            return await self.handle_generation_request(req)

        @app.get("/v1/models")
        async def list_models():
            return {"object": "list", "data": [{"id": self.model_path, "object": "model",
                                               "created": int(self.started), "owned_by": "llmctl"}]}

        @app.get("/health")
        async def health():
            return {"status": "healthy" if self.loop_thread.error is None else "degraded",
                    "active_requests": len(self.active_requests), "pending_requests": self.scheduler.num_waiting,
                    "running_requests": self.scheduler.num_running, "kv_free_blocks": self.engine.kv.free_blocks,
                    "kv_total_blocks": self.engine.kv.num_blocks}

        @app.get("/metrics")
        async def metrics():
            try:
                from prometheus_client import generate_latest
                return PlainTextResponse(generate_latest().decode())
            except Exception:  # noqa: BLE001
                return PlainTextResponse("")

    # ---- request handling ----------------------------------------------------------------------------
    def _record(self, seq: Sequence) -> None:
        try:
            from ..metrics.observability import get_observability_manager
            lat = time.time() - seq.created
            get_observability_manager().record_inference_request(
                latency=lat, tokens=len(seq.generated), ttft=(seq.first_token_time or time.time()) - seq.created,
                queue_depth=self.scheduler.num_waiting, error=seq.finish_reason == "error")
            from ..metrics.health import get_health_manager
            get_health_manager().update_inference_metrics(latency=lat, error=seq.finish_reason == "error",
                                                          queue_size=self.scheduler.num_waiting)
        except Exception:  # noqa: BLE001
            pass

    def _submit(self, req: GenerationRequest, on_token=None):
        ids = self.engine.tokenizer.encode(req.prompt)
        seq = self.engine.add_request(ids, req.max_tokens, req.temperature, req.top_p, req.top_k, on_token=on_token)
        rid = str(uuid.uuid4())
        self.active_requests[rid] = RequestState(rid, seq)
        self.loop_thread.wake.set()
        return rid, seq, len(ids)

    def _response(self, rid: str, seq: Sequence, n_prompt: int) -> GenerationResponse:
        gen = [t for t in seq.generated if t != getattr(self.engine.tokenizer, "eos_token_id", -1)]
        text = self.engine.tokenizer.decode(gen)
        return GenerationResponse(id=rid, text=text, finish_reason=seq.finish_reason or "error",
                                  usage={"prompt_tokens": n_prompt, "completion_tokens": len(seq.generated),
                                         "total_tokens": n_prompt + len(seq.generated)})

    async def handle_generation_request(self, req: GenerationRequest) -> GenerationResponse:
        loop = asyncio.get_running_loop()
        # off the event loop: under tensor-parallel (lockstep) serving admission waits for the next engine step
        rid, seq, n_prompt = await loop.run_in_executor(None, self._submit, req)
        try:
            await loop.run_in_executor(None, seq.done_event.wait)
            self.completed += 1
            if seq.finish_reason == "error":
                self.errors += 1
            return self._response(rid, seq, n_prompt)
        finally:
            if not seq.finished:
                seq.cancel_requested = True
            self.active_requests.pop(rid, None)

    async def _stream(self, req: GenerationRequest):
        loop = asyncio.get_running_loop()
        q: asyncio.Queue = asyncio.Queue()

        def on_token(seq, tok):
            loop.call_soon_threadsafe(q.put_nowait, tok)

        rid, seq, n_prompt = await loop.run_in_executor(None, self._submit, req, on_token)
        try:
            sent = 0
            while True:
                if seq.finished and q.empty():
                    break
                try:
                    tok = await asyncio.wait_for(q.get(), timeout=0.1)
                except asyncio.TimeoutError:
                    continue
                sent += 1
                piece = self.engine.tokenizer.decode([tok])
                yield f"data: {json.dumps({'id': rid, 'text': piece, 'index': sent - 1})}\n\n"
            yield f"data: {json.dumps({'id': rid, 'finish_reason': seq.finish_reason, 'usage': self._response(rid, seq, n_prompt).usage})}\n\n"
            yield "data: [DONE]\n\n"
            self.completed += 1
        finally:
            if not seq.finished:
                seq.cancel_requested = True        # client disconnected: stop generating for it
            self.active_requests.pop(rid, None)

    # ---- lifecycle -----------------------------------------------------------------------------------------
    def start_background(self) -> None:
        if not self.loop_thread.is_alive():
            self.loop_thread.start()
        try:
            # error rate / latency / queue depth monitors (fed by _record) raise alerts in the server log
            import logging
            from ..metrics.health import get_health_manager
            hm = get_health_manager()
            if not getattr(hm, "_llmctl_serve_wired", False):
                log = logging.getLogger("llmctl.serve")
                hm.add_alert_callback(lambda r: log.warning("health %s: %s - %s", r.status.value.upper(), r.component, r.message)
                                      if r.component != "training" else None)
                hm.start_monitoring()
                hm._llmctl_serve_wired = True
        except Exception:  # noqa: BLE001 - monitoring is best effort
            pass

    async def start_server(self):
        import uvicorn
        self.start_background()
        config = uvicorn.Config(self.app, host=self.host, port=self.port, log_level="info")
        await uvicorn.Server(config).serve()

    def shutdown(self):
        self.loop_thread.stop()
        if hasattr(self.engine, "stop"):                  # lockstep leader: release the follower ranks
            self.loop_thread.join(timeout=5.0)
            self.engine.stop()


def create_inference_server(model_path: str, **kwargs) -> InferenceServer:
    """Factory kept from reference llmctl/serve/server.py:409-411."""
    return InferenceServer(model_path, **kwargs)


# names re-exported for parity with the reference's module
KVCacheManager = None
DynamicBatchScheduler = None


def _late_bind():
    global KVCacheManager, DynamicBatchScheduler
    from .native_host import BatchScheduler, KVBlockManager
    KVCacheManager, DynamicBatchScheduler = KVBlockManager, BatchScheduler


_late_bind()
