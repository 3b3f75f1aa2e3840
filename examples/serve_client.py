#!/usr/bin/env python
"""Start the HTTP server on a free port in this process, send concurrent completion requests (one of them streamed)
and print per-request latency — the counterpart of the reference's test_inference.py, which needed a server started
by hand and could only ever get one token back (reference llmctl/serve/server.py:372-386).

    python examples/serve_client.py [--artifact DIR_OR_PRESET] [--requests 4] [--max-tokens 16]
"""
import argparse
import asyncio
import json
import socket
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


async def main_async(a):
    import aiohttp
    import uvicorn
    from llmctl_b200.serve.server import create_inference_server
    server = create_inference_server(a.artifact, max_batch_size=a.requests, max_batch_tokens=4096)
    server.start_background()                                   # the engine loop thread (start_server() does this too)
    port = free_port()
    uv = uvicorn.Server(uvicorn.Config(server.app, host="127.0.0.1", port=port, log_level="warning"))
    task = asyncio.create_task(uv.serve())
    base = f"http://127.0.0.1:{port}"
    async with aiohttp.ClientSession() as s:
        for _ in range(100):                                   # wait for the listener
            try:
                async with s.get(base + "/health") as r:
                    if r.status == 200:
                        break
            except aiohttp.ClientError:
                await asyncio.sleep(0.05)
        async with s.get(base + "/v1/models") as r:
            print("models:", json.dumps(await r.json())[:200])

        async def one(i):
            t0 = time.time()
            body = {"prompt": f"request {i}: the quick brown fox", "max_tokens": a.max_tokens, "temperature": 0.7, "top_p": 0.9}
            async with s.post(base + "/v1/completions", json=body) as r:
                d = await r.json()
            print(f"request {i}: {d['usage']['completion_tokens']} tokens, finish={d['finish_reason']}, "
                  f"{1e3 * (time.time() - t0):.0f} ms, text={d['text'][:40]!r}")
            return d["usage"]["completion_tokens"]

        async def streamed():
            n = 0
            body = {"prompt": "streamed request", "max_tokens": a.max_tokens, "temperature": 0.0, "stream": True}
            async with s.post(base + "/v1/completions", json=body) as r:
                async for line in r.content:
                    if line.startswith(b"data: ") and b"[DONE]" not in line:
                        n += 1
            print(f"streamed request: {n} server-sent events")
            return n

        counts = await asyncio.gather(*[one(i) for i in range(a.requests)], streamed())
        async with s.get(base + "/health") as r:
            print("health:", json.dumps(await r.json())[:300])
    uv.should_exit = True
    await task
    server.shutdown()
    assert all(c >= 1 for c in counts), counts


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--artifact", default="gpt-tiny")
    ap.add_argument("--requests", type=int, default=4)
    ap.add_argument("--max-tokens", type=int, default=16)
    asyncio.run(main_async(ap.parse_args()))
