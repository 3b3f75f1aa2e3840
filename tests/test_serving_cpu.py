"""Host-side serving logic on CPU: C++ KV block manager, C++ continuous-batching scheduler (the reference's
budget semantics minus head-of-line blocking), sampling semantics vs the reference pipeline, and the engine
end to end (prefill + paged decode) against a plain full-recompute forward."""
import asyncio
import json

import pytest
import torch


def test_kv_block_manager():
    from llmctl_b200.serve.native_host import KVBlockManager
    kv = KVBlockManager(num_blocks=8, block_size=4, watermark=0.0)
    assert kv.allocate(1, 5) and kv.block_table(1) == [0, 1] and kv.free_blocks == 6
    assert kv.allocate(1, 8) and len(kv.block_table(1)) == 2          # still fits in two blocks
    assert kv.allocate(1, 9) and len(kv.block_table(1)) == 3
    assert not kv.allocate(2, 4 * 6)                                   # needs 6 blocks, only 5 free -> unchanged
    assert kv.free_blocks == 5 and kv.num_tokens(2) == -1
    assert kv.allocate(2, 4 * 5) and kv.free_blocks == 0 and not kv.can_allocate(1)
    assert kv.free(1) == 3 and kv.free_blocks == 3
    # fork shares full blocks, copies the partial one
    kv2 = KVBlockManager(8, 4, 0.0)
    kv2.allocate(10, 6)
    cow = kv2.fork(10, 11)
    assert cow is not None and cow[0] == kv2.block_table(10)[1] and cow[1] == kv2.block_table(11)[1]
    assert kv2.block_table(10)[0] == kv2.block_table(11)[0]
    assert kv2.free(10) == 1 and kv2.free(11) == 2 and kv2.free_blocks == 8


def test_scheduler_budgets_no_hol_blocking():
    from llmctl_b200.serve.native_host import DECODE, PREFILL, BatchScheduler
    s = BatchScheduler(max_batch_size=2, max_batch_tokens=100)
    assert not s.add_request(99, 90, 20)            # can never fit: rejected instead of blocking the queue forever
    assert s.add_request(1, 40, 20)                 # reserves 60
    assert s.add_request(2, 50, 20)                 # 70: does not fit next to #1
    assert s.add_request(3, 10, 10)                 # 20: fits -> skips ahead of #2 (no head-of-line blocking)
    plan = s.step()
    assert [(i, p) for i, p, _ in plan] == [(1, PREFILL), (3, PREFILL)] and plan[0][2] == 40
    assert [(i, p, n) for i, p, n in s.step()] == [(1, DECODE, 1), (3, DECODE, 1)]
    s.finish(1)
    plan = s.step()                                 # continuous batching: #2 joins while #3 keeps decoding
    assert (2, PREFILL, 50) in plan and (3, DECODE, 1) in plan
    assert s.num_waiting == 0 and s.num_running == 2


def test_scheduler_simple_policy_and_preemption():
    from llmctl_b200.serve.native_host import PREFILL, BatchScheduler, KVBlockManager
    s = BatchScheduler(2, 1000, "simple")
    for i in (1, 2, 3):
        s.add_request(i, 10, 5)
    assert {i for i, _, _ in s.step()} == {1, 2}
    s.finish(1)
    assert {i for i, _, _ in s.step()} == {2}       # static batching: nothing joins until the batch drains
    s.finish(2)
    assert {i for i, _, _ in s.step()} == {3}
    kv = KVBlockManager(4, 4, 0.0)                   # 16 token slots in total
    v = BatchScheduler(4, 1000, "vllm-like", kv)
    v.add_request(1, 7, 8); v.add_request(2, 7, 8)
    assert {i for i, _, _ in v.step()} == {1, 2} and kv.free_blocks == 0
    assert v.on_token(1)                             # 8 tokens: still 2 blocks
    assert v.on_token(1) is True                     # 9 tokens: needs a 3rd block -> newest sequence (#2) is preempted
    assert v.pop_preempted() == [2] and v.num_running == 1 and v.num_waiting == 1
    v.finish(1)
    assert (2, PREFILL, 7) in v.step()               # re-admitted and re-prefilled


def test_sampling_semantics_match_reference_pipeline():
    from llmctl_b200.ops.serving import sample_reference
    torch.manual_seed(0)
    logits = torch.tensor([2.0, 1.0, 0.5, 0.1, -1.0, -3.0])
    assert sample_reference(logits, 0.0, -1, 1.0) == 0                       # greedy
    g = torch.Generator().manual_seed(1)
    draws = {sample_reference(logits, 1.0, 2, 1.0, g) for _ in range(200)}
    assert draws == {0, 1}                                                   # top-k keeps the 2 largest
    p = torch.softmax(logits, -1)
    # nucleus: tokens are dropped only when the cumulative probability BEFORE them exceeds p -> first crossing token kept
    top_p = float(p[0]) + 1e-4
    draws = {sample_reference(logits, 1.0, -1, top_p, g) for _ in range(300)}
    assert draws == {0, 1}
    draws = {sample_reference(logits, 1.0, -1, float(p[0]) - 1e-4, g) for _ in range(100)}
    assert draws == {0}


def _tiny_engine(**kw):
    from llmctl_b200.comms import mesh
    from llmctl_b200.serve.engine import InferenceEngine
    mesh.reset_state()
    torch.manual_seed(0)
    return InferenceEngine("gpt-tiny", device="cpu", kv_block_size=4, num_kv_blocks=64, **kw)


def test_engine_paged_decode_matches_full_recompute():
    eng = _tiny_engine(max_batch_size=4, max_batch_tokens=512)
    prompts = [[5, 9, 200, 31, 77], [8, 8, 100], [300, 2, 3, 4, 5, 6, 7, 401, 9]]
    outs = eng.generate(prompts, max_tokens=6, temperature=0.0)
    assert all(len(o) == 6 for o in outs)
    model = eng.model
    for p, o in zip(prompts, outs):                   # oracle: greedy decoding by recomputing the whole sequence
        ids = list(p)
        for t in o:
            logits = model(torch.tensor([ids])).float()[-1, :eng.cfg.vocab_size]
            assert int(logits.argmax()) == t
            ids.append(t)
    assert eng.kv.free_blocks == 64 and not eng.has_work()      # every block returned


def test_engine_continuous_batching_under_pressure():
    eng = _tiny_engine(max_batch_size=2, max_batch_tokens=64, scheduler="vllm-like")
    seqs = [eng.add_request([i + 3, i + 4, i + 5], max_tokens=5, temperature=0.0) for i in range(5)]
    steps = 0
    while eng.has_work():
        eng.step(); steps += 1
        assert steps < 200
    assert all(s.finished and s.finish_reason == "length" and len(s.generated) == 5 for s in seqs)
    too_big = eng.add_request(list(range(3, 60)), max_tokens=50)
    assert too_big.finished and too_big.finish_reason == "error"


def test_engine_preemption_under_kv_pressure_matches_full_recompute():
    """Eight KV blocks of four tokens for four requests that each need four: growth evicts running sequences in the
    middle of a step (both from the prefill phase and inside the batched decode).  Evicted sequences are re-prefilled
    over prompt + generated tokens and every output must still equal greedy decoding by full recompute."""
    from llmctl_b200.comms import mesh
    from llmctl_b200.serve.engine import InferenceEngine
    mesh.reset_state()
    torch.manual_seed(0)
    eng = InferenceEngine("gpt-tiny", device="cpu", kv_block_size=4, num_kv_blocks=8, max_batch_size=4,
                          max_batch_tokens=512, scheduler="vllm-like")
    prompts = [[5, 9, 200, 31], [8, 8, 100], [300, 2, 3, 4, 5], [17, 18]]
    outs = eng.generate(prompts, max_tokens=12, temperature=0.0)
    assert eng.stats["preemptions"] >= 1, "the test did not create KV pressure"
    assert all(len(o) == 12 for o in outs), [len(o) for o in outs]
    for p, o in zip(prompts, outs):
        ids = list(p)
        for t in o:
            logits = eng.model(torch.tensor([ids])).float()[-1, :eng.cfg.vocab_size]
            assert int(logits.argmax()) == t
            ids.append(t)
    assert eng.kv.free_blocks == 8 and not eng.has_work()


def test_request_larger_than_the_watermarked_pool_is_not_starved():
    """A request that needs more blocks than (total - watermark) is admitted once the engine is empty instead of
    waiting forever; generate() terminates either way."""
    from llmctl_b200.comms import mesh
    from llmctl_b200.serve.engine import InferenceEngine
    mesh.reset_state()
    torch.manual_seed(0)
    eng = InferenceEngine("gpt-tiny", device="cpu", kv_block_size=4, num_kv_blocks=8, max_batch_size=2,
                          max_batch_tokens=512, scheduler="vllm-like")
    out = eng.generate([list(range(3, 3 + 26))], max_tokens=5, temperature=0.0)      # 31 tokens = 8 blocks = whole pool
    assert len(out[0]) == 5 and not eng.has_work()


def test_scheduler_plugin_orders_admission():
    """A registered scheduler plugin (``schedule(requests) -> ordered requests``) decides which waiting request the
    native continuous-batching core sees first; budgets and KV accounting stay native."""
    from llmctl_b200.plugins.registry import get_registry

    class ShortestFirst:
        def schedule(self, requests):
            return sorted(requests, key=lambda r: r.prompt_len)

    get_registry().register("schedulers", "shortest-first", ShortestFirst, overwrite=True)
    eng = _tiny_engine(max_batch_size=1, max_batch_tokens=256, scheduler="shortest-first")
    long_ = eng.add_request(list(range(3, 33)), max_tokens=3, temperature=0.0)
    short = eng.add_request([5, 6, 7], max_tokens=3, temperature=0.0)
    first_done = None
    steps = 0
    while eng.has_work():
        for s in eng.step():
            first_done = first_done or s
        steps += 1
        assert steps < 100
    assert first_done is short, "the plugin's order (shortest prompt first) was not followed"
    assert long_.finished and len(long_.generated) == 3 and len(short.generated) == 3
    assert eng.kv.free_blocks == 64


def test_http_api_contract():
    from fastapi.testclient import TestClient
    from llmctl_b200.serve.server import InferenceServer
    eng = _tiny_engine(max_batch_size=4, max_batch_tokens=512)
    srv = InferenceServer("gpt-tiny", engine=eng, max_concurrent_requests=2)
    srv.start_background()
    try:
        c = TestClient(srv.app)
        h = c.get("/health").json()
        assert h["status"] == "healthy" and h["active_requests"] == 0 and "pending_requests" in h
        m = c.get("/v1/models").json()
        assert m["object"] == "list" and m["data"][0]["owned_by"] == "llmctl" and m["data"][0]["object"] == "model"
        r = c.post("/v1/completions", json={"prompt": "hello", "max_tokens": 7, "temperature": 0.0})
        assert r.status_code == 200
        body = r.json()
        assert set(body) == {"id", "text", "finish_reason", "usage"} and body["finish_reason"] in ("stop", "length")
        assert body["usage"]["prompt_tokens"] == 5 and 1 <= body["usage"]["completion_tokens"] <= 7
        assert body["usage"]["total_tokens"] == body["usage"]["prompt_tokens"] + body["usage"]["completion_tokens"]
        with c.stream("POST", "/v1/completions", json={"prompt": "hi", "max_tokens": 4, "temperature": 0.0, "stream": True}) as s:
            lines = [l for l in s.iter_lines() if l.startswith("data: ")]
        assert lines[-1] == "data: [DONE]" and len(lines) >= 3
        assert c.post("/v1/completions", json={"prompt": "x", "max_tokens": 0}).status_code == 422
        import time as _t
        _t.sleep(0.2)                                    # the engine loop records finished requests
        m = c.get("/metrics").text
        assert 'llmctl_inference_requests_total{status="ok"}' in m and "llmctl_inference_latency_seconds_count" in m
    finally:
        srv.shutdown()


def test_cancelled_sequence_frees_its_blocks_early():
    eng = _tiny_engine(max_batch_size=2, max_batch_tokens=256)
    a = eng.add_request([5, 6, 7], max_tokens=50, temperature=0.0)
    b = eng.add_request([8, 9], max_tokens=6, temperature=0.0)
    eng.step(); eng.step()
    a.cancel_requested = True
    steps = 0
    while eng.has_work():
        eng.step(); steps += 1
    assert a.finished and a.finish_reason == "cancelled" and len(a.generated) < 10
    assert b.finish_reason == "length" and len(b.generated) == 6
    assert eng.kv.free_blocks == 64 and steps < 12


def test_tensor_parallel_http_serving_lockstep(tmp_path):
    """`llmctl serve --tensor-parallel 2` on CPU/gloo: two processes, each with half of every weight matrix; rank 0
    answers HTTP, the follower mirrors its engine step by step (serve/lockstep.py).  Concurrent completions finish with
    the requested number of tokens, a greedy request is reproducible, /health reports the shared KV pool."""
    import os
    import subprocess
    import sys
    import time
    import urllib.request
    from pathlib import Path
    from dist_utils import free_port
    root = Path(__file__).resolve().parent.parent
    port = free_port()
    worker = root / "distributed-llm-training-and-inference-system_b200" / "serve" / "tp_worker.py"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), str(worker), "--artifact", "gpt-tiny", "--host", "127.0.0.1", "--port", str(port),
           "--device", "cpu", "--max-batch-size", "4", "--num-kv-blocks", "256"]
    proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                            env=dict(os.environ, PYTHONPATH=str(root), OMP_NUM_THREADS="2"), start_new_session=True)

    def post(payload, timeout=120):
        req = urllib.request.Request(f"http://127.0.0.1:{port}/v1/completions", data=json.dumps(payload).encode(),
                                     headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=timeout) as r:
            return json.loads(r.read())
    try:
        deadline = time.time() + 120
        while True:
            try:
                with urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=2) as r:
                    health = json.loads(r.read())
                break
            except Exception:  # noqa: BLE001 - not up yet
                assert proc.poll() is None, proc.stdout.read()[-3000:]
                assert time.time() < deadline, "server did not come up"
                time.sleep(0.5)
        assert health["status"] == "healthy" and health["kv_total_blocks"] == 256
        import concurrent.futures as cf
        with cf.ThreadPoolExecutor(4) as ex:
            futs = [ex.submit(post, {"prompt": f"hello {i}", "max_tokens": 6 + i, "temperature": 0.0}) for i in range(4)]
            outs = [f.result() for f in futs]
        for i, o in enumerate(outs):
            assert o["finish_reason"] in ("length", "stop") and o["usage"]["completion_tokens"] <= 6 + i
            assert o["usage"]["completion_tokens"] >= 1
        a = post({"prompt": "the same prompt", "max_tokens": 8, "temperature": 0.0})
        b = post({"prompt": "the same prompt", "max_tokens": 8, "temperature": 0.0})
        assert a["text"] == b["text"] and a["usage"] == b["usage"]
        # a streaming client that goes away after the first event: the cancellation must reach the follower too, or the
        # two ranks would disagree about the batch from the next step on and the following request could not complete
        import socket
        body = json.dumps({"prompt": "abandoned", "max_tokens": 200, "temperature": 0.0, "stream": True}).encode()
        sk = socket.create_connection(("127.0.0.1", port), timeout=30)
        sk.sendall(b"POST /v1/completions HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: "
                   + str(len(body)).encode() + b"\r\n\r\n" + body)
        got = b""
        while b"data:" not in got:
            got += sk.recv(4096)
        sk.close()
        c = post({"prompt": "after the cancel", "max_tokens": 5, "temperature": 0.0})
        assert c["usage"]["completion_tokens"] >= 1
        deadline = time.time() + 30
        while True:
            with urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=5) as r:
                h2 = json.loads(r.read())
            if h2["kv_free_blocks"] == 256 and h2["active_requests"] == 0:          # every block returned, also the abandoned one's
                break
            assert time.time() < deadline, h2
            time.sleep(0.2)
    finally:
        import signal
        try:
            os.killpg(proc.pid, signal.SIGTERM)
        except ProcessLookupError:
            pass
        try:
            proc.wait(timeout=20)
        except subprocess.TimeoutExpired:
            os.killpg(proc.pid, signal.SIGKILL)


def test_bench_e2e_tensor_parallel_cli_on_gloo():
    """`llmctl bench e2e --tp 2` under torchrun on CPU/gloo: the serving benchmark's tensor-parallel path (mesh set-up,
    identical admission on both ranks, orderly teardown) — rank 0 prints the result."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    from dist_utils import free_port
    root = Path(__file__).resolve().parent.parent
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), "-m", "llmctl_b200.cli.main", "bench", "e2e", "--tp", "2", "--model", "gpt-tiny",
           "--prompt-length", "16", "--gen-length", "4", "--num-requests", "2", "--warmup-runs", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=str(root),
                       env=dict(os.environ, PYTHONPATH=str(root), OMP_NUM_THREADS="2"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = r.stdout
    res = json.loads(out[out.index("{"):out.rindex("}") + 1])
    assert res["tensor_parallel"] == 2 and res["steps"] == 4 and res["decode_tokens_per_s"] > 0
