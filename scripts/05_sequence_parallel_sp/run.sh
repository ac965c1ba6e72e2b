#!/usr/bin/env bash
# Launch the recipes of this directory on one NVSwitch box, one after the other (replaces the reference's PBS + mpiexec
# templates, which exist only for the strategies the reference implements as code; these two are docs-only there).
set -euo pipefail
NPROC=${NPROC:-$(python -c 'import torch; print(max(2, torch.cuda.device_count()))')}
HERE=$(cd "$(dirname "$0")" && pwd)
for s in "$HERE"/0*.py; do
  echo "=== $(basename "$s") on $NPROC processes ==="
  "$HERE/../launch.sh" "$NPROC" "$s" "$@"
done
