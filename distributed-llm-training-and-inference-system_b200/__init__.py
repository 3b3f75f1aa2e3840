"""llmctl-b200: a Blackwell-native (sm_100a) LLM training and serving framework with the CLI surface,
file formats and checkpoint layout of ``llmctl`` (ambicuity/Distributed-LLM-Training-and-Inference-System).

Layer map (mirrors reference README.md:14-25, all ten layers are real here):

    cli/        typer app, 13 command groups           (reference llmctl/cli/)
    runtime/    launchers, worker entry, train engine  (reference llmctl/runtime/)
    partition/  planner + sharding maps                (reference: empty package; planner in cli/commands/plan.py)
    exec/ ops/  sm_100a kernels + autograd wrappers    (reference: empty package)
    comms/      symmetric memory, fused collectives    (reference: empty package)
    parallel/   DP / ZeRO / TP / SP / PP               (reference: plan integers only)
    io/         datasets, checkpointing, export        (reference: empty package)
    serve/      HTTP server, scheduler, paged KV       (reference llmctl/serve/)
    metrics/    observability, health, roofline        (reference llmctl/metrics/)
    config/     pydantic schemas + presets             (reference: empty package)
    plugins/    registry + auto-tuner                  (reference llmctl/plugins/)
"""

__version__ = "0.1.0"
__description__ = "Distributed LLM Training and Inference System (B200-native)"
