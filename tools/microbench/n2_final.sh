#!/bin/bash
# 2-GPU closing check of the final tree: strategy parity + headline bench
N=2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=60
timeout 400 $TR tools/parallel_check.py > gpurun_out/parallel_check_n2_final.out 2>&1; echo "parallel_check rc=$? pass=$(grep -c PASS gpurun_out/parallel_check_n2_final.out)"; grep "FAIL\|failures" gpurun_out/parallel_check_n2_final.out | head -5
timeout 170 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n2_final.out 2> gpurun_out/bench_n2_final.err; echo "bench rc=$?"; grep '^{' gpurun_out/bench_n2_final.out | cut -c1-200
