#!/bin/bash
# 8-GPU closing run: fused-TP hybrid (after the split-handshake fix), collective checks and the full sweep with the registered arm
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=45
timeout 170 $TR bench/hybrid_bench.py --tp 4 --fused --steps 6 --warmup 3 > gpurun_out/hybrid_fused_n8.out 2> gpurun_out/hybrid_fused_n8.err; echo "hybrid_fused rc=$?"; grep '^{' gpurun_out/hybrid_fused_n8.out | cut -c1-260; grep -m3 watchdog gpurun_out/hybrid_fused_n8.out gpurun_out/hybrid_fused_n8.err
timeout 170 $TR tools/comm_check.py > gpurun_out/comm_check_n8.out 2>&1; echo "comm_check rc=$? pass=$(grep -c PASS gpurun_out/comm_check_n8.out)"; grep "FAIL\|failures\|TIME" gpurun_out/comm_check_n8.out | head -14
tools/microbench/fsdp_ctas_sweep.sh 8 "64 128"
rm -f gpurun_out/comm_bench_8gpu.csv
timeout 330 $TR bench/comm_bench.py --impl both --ops broadcast all_reduce all_gather reduce_scatter --benchmark_runs 10 --warmup_runs 3 --output_file gpurun_out/comm_bench_8gpu.csv > gpurun_out/comm_bench_n8.out 2> gpurun_out/comm_bench_n8.err; echo "comm_bench rc=$? rows=$(grep -c , gpurun_out/comm_bench_8gpu.csv)"
