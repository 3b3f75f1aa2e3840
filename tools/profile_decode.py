"""Decode-step kernel breakdown: prefill N prompts, then run a few eager decode steps between
cudaProfilerStart/Stop so that `ncu --profile-from-start off --metrics gpu__time_duration.sum` lists them.
    ncu --profile-from-start off --metrics gpu__time_duration.sum --csv --log-file out.csv python tools/profile_decode.py
Without ncu it prints per-phase wall times (prefill per request, decode per step) for graph and eager modes."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.serve.engine import InferenceEngine  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "gpt-7b"
graphs = (sys.argv[2] == "graph") if len(sys.argv) > 2 else False
nreq, plen = 8, 2048
eng = InferenceEngine(model, max_batch_size=nreq, max_batch_tokens=nreq * (plen + 256), cuda_graphs=graphs)
eng.tokenizer.eos_token_id = -1
g = torch.Generator().manual_seed(0)
for rnd in range(2):
    seqs = [eng.add_request(torch.randint(3, 30000, (plen,), generator=g).tolist(), 24, temperature=0.0) for _ in range(nreq)]
    t0 = time.perf_counter()
    eng.step()                                  # prefill of all admitted requests (+ nothing to decode yet)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for _ in range(4):
        eng.step()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    if rnd == 1:
        torch.cuda.profiler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(3):
        eng.step()
    ev1.record()
    torch.cuda.synchronize()
    if rnd == 1:
        torch.cuda.profiler.stop()
    t3 = time.perf_counter()
    print(f"round {rnd} graphs={graphs}: prefill step {1e3 * (t1 - t0):.1f} ms, first 4 decode steps {1e3 * (t2 - t1) / 4:.2f} ms/step, "
          f"next 3 {1e3 * (t3 - t2) / 3:.2f} ms/step (device {ev0.elapsed_time(ev1) / 3:.2f} ms/step), stats {eng.stats}")
    while eng.has_work():
        eng.step()
