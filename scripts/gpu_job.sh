#!/usr/bin/env bash
# One parameterised runner for gpurun sessions: `scripts/gpu_job.sh <name> -- <command...>` runs the command from the
# repo root with its stdout/stderr captured under gpurun_out/<name>.log and the exit code in gpurun_out/<name>.rc.
# Several jobs can be chained in one gpurun call:  gpurun -- 'scripts/gpu_job.sh a -- cmd1; scripts/gpu_job.sh b -- cmd2'
set -u
name="$1"; shift; [ "$1" = "--" ] && shift
mkdir -p gpurun_out
export PYTHONPATH="$(pwd)${PYTHONPATH:+:$PYTHONPATH}"
( "$@" ) > "gpurun_out/${name}.log" 2>&1
rc=$?
echo "$rc" > "gpurun_out/${name}.rc"
echo "[gpu_job] ${name}: rc=${rc}"; tail -n 6 "gpurun_out/${name}.log"
exit 0
