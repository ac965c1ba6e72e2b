#!/usr/bin/env bash
# Launch the recipes of this directory on one NVSwitch box (replaces the reference's PBS + mpiexec templates).
set -euo pipefail
NPROC=${NPROC:-$(python -c 'import torch; print(max(1, torch.cuda.device_count()))')}
HERE=$(cd "$(dirname "$0")" && pwd)
exec "$HERE/../launch.sh" "$NPROC" "$HERE/01_fsdp_tp_hybrid.py"  "$@"
