#!/bin/bash
# N-GPU experiment: registered-path correctness, then reduce-scatter kernel variants and registered AG / broadcast
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=60
timeout 200 $TR tools/comm_check.py > gpurun_out/comm_check_reg_n$N.out 2>&1; echo "comm_check rc=$?"; grep -c PASS gpurun_out/comm_check_reg_n$N.out; grep "FAIL\|registered\|failures" gpurun_out/comm_check_reg_n$N.out | head -20
for t in 0; do
  B200DIST_NVLS_RS_TUNE=$t timeout 120 $TR bench/comm_bench.py --impl b200dist --ops reduce_scatter --tensor_sizes 16777216 67108864 268435456 --benchmark_runs 10 --warmup_runs 3 --output_file gpurun_out/rs_tune${t}_n$N.csv > gpurun_out/rs_tune${t}_n$N.out 2>&1
  echo "tune $t rc=$?"; grep "^b200dist" gpurun_out/rs_tune${t}_n$N.out | cut -d, -f1,2,4,10,12,14
done
timeout 200 $TR bench/comm_bench.py --impl both --ops all_gather broadcast reduce_scatter --tensor_sizes 4194304 67108864 268435456 --benchmark_runs 10 --warmup_runs 3 --output_file gpurun_out/reg_n$N.csv > gpurun_out/reg_n$N.out 2>&1
echo "reg rc=$?"; grep "^b200dist\|^nccl" gpurun_out/reg_n$N.out | cut -d, -f1,2,4,10,12,14
