#!/usr/bin/env bash
# torchrun on this node with N ranks on a free port:  scripts/torchrun_local.sh N script.py [args...]
n="$1"; shift
port=$(python - <<'PY'
import socket
s = socket.socket(); s.bind(("127.0.0.1", 0)); print(s.getsockname()[1])
PY
)
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"
