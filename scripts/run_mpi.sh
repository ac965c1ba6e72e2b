#!/usr/bin/env bash
# MPI launch of the ResNet benchmark - counterpart of the reference's scripts/run_mpi.sh (mpiexec -np 8 ./main.py
# --backend nccl, three repetitions, :99-127).  One task per GPU; b200dist.utils.init_distributed picks the rank
# variables of Open MPI / MPICH-PMI / PALS up by itself and rank 0 publishes the rendezvous address, so no
# Slingshot / libfabric / OFI-plugin environment is needed on one NVSwitch box.
#   scripts/run_mpi.sh [nproc] [extra main.py flags...]
set -euo pipefail
cd "$(dirname "$0")/.."
NP=${1:-8}; shift || true
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1} MASTER_PORT=${MASTER_PORT:-29500}
MPIEXEC=${MPIEXEC:-mpiexec}
command -v "$MPIEXEC" >/dev/null || { echo "no $MPIEXEC on PATH; use scripts/launch.sh $NP scripts/main.py ..." >&2; exit 2; }
for rep in 1 2 3; do
  echo "--- repetition $rep ---"
  "$MPIEXEC" -n "$NP" python scripts/main.py --backend nccl --arch resnet50 --use_syn --batch_size 256 "$@"
done
