#!/bin/bash
N=2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=60
timeout 200 $TR tools/comm_check.py > gpurun_out/comm_check_n2_split.out 2>&1; echo "comm_check rc=$? pass=$(grep -c PASS gpurun_out/comm_check_n2_split.out)"; grep "FAIL\|split\|failures" gpurun_out/comm_check_n2_split.out | head -12
timeout 400 $TR tools/parallel_check.py > gpurun_out/parallel_check_n2_split.out 2>&1; echo "parallel_check rc=$?"; grep -c PASS gpurun_out/parallel_check_n2_split.out; grep "FAIL\|TIME\|failures" gpurun_out/parallel_check_n2_split.out | head -12
tools/microbench/fsdp_ctas_sweep.sh 2 "32 64 128"
tools/multigpu_suite.sh 2 resnet
