#!/usr/bin/env bash
# Single-node launcher for every recipe:  scripts/launch.sh <nproc> <script.py> [args...]
# Replaces the reference's PBS/mpiexec templates (scripts/*/*.sh): on one NVSwitch box none of the
# Slingshot / libfabric / MPICH variables apply, so the only setup is the rendezvous address.
set -euo pipefail
NPROC=${1:?usage: launch.sh <nproc> <script.py> [args...]}; shift
SCRIPT=${1:?script}; shift
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}
export MASTER_PORT=${MASTER_PORT:-29500}
export NCCL_DEBUG=${NCCL_DEBUG:-VERSION}
if [ "$NPROC" -le 1 ]; then exec python "$SCRIPT" "$@"; fi
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NPROC" --master-addr "$MASTER_ADDR" \
     --master-port "$MASTER_PORT" "$SCRIPT" "$@"
