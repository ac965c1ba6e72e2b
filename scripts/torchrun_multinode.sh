#!/usr/bin/env bash
# Multi-node torchrun template - counterpart of the reference's scripts/torchrun_multigpu_pbs.sh (mpiexec -n NNODES --ppn 1
# torchrun --nnodes=N --nproc-per-node=4 --rdzv-backend=c10d --rdzv-endpoint=head:port main.py, :152-161).  The peer-memory
# data plane of this framework spans ONE NVSwitch domain: across nodes use it for the intra-node dimension of a mesh (TP /
# FSDP inside the box) and let torch.distributed (NCCL) carry the inter-node dimension (DP / PP).
#   NNODES=2 HEAD=node0 scripts/torchrun_multinode.sh scripts/main.py --arch resnet50 --use_syn
set -euo pipefail
NNODES=${NNODES:-1}; HEAD=${HEAD:-127.0.0.1}; PORT=${PORT:-29500}; GPUS=${GPUS:-8}
exec python -m torch.distributed.run --nnodes="$NNODES" --nproc-per-node="$GPUS" --rdzv-backend=c10d \
     --rdzv-endpoint="$HEAD:$PORT" "$@"
