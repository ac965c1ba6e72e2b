#!/bin/bash
# 8-GPU retest after the 256-thread split-mode bulk kernels: fused-TP hybrid, collective checks, headline bench
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=40
timeout 160 $TR bench/hybrid_bench.py --tp 4 --fused --steps 6 --warmup 3 > gpurun_out/hybrid_fused_n8.out 2> gpurun_out/hybrid_fused_n8.err; echo "hybrid_fused rc=$?"; grep '^{' gpurun_out/hybrid_fused_n8.out | cut -c1-260; grep -h watchdog gpurun_out/hybrid_fused_n8.out | sed 's/block [0-9]*/block N/' | sort | uniq -c | sort -rn | head -4
timeout 150 $TR tools/comm_check.py > gpurun_out/comm_check_n8.out 2>&1; echo "comm_check rc=$? pass=$(grep -c PASS gpurun_out/comm_check_n8.out)"; grep "FAIL\|failures" gpurun_out/comm_check_n8.out | head -5
timeout 170 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n8_split256.out 2> gpurun_out/bench_n8_split256.err; echo "bench rc=$?"; grep '^{' gpurun_out/bench_n8_split256.out | cut -c1-200
