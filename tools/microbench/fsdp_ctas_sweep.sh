#!/bin/bash
# N-GPU sweep of the CTA cap of the FSDP collective kernels (reduce-scatter + AdamW through the switch runs beside the
# backward GEMMs / attention): tools/microbench/fsdp_ctas_sweep.sh N "32 64 128"
N=$1; CTAS=${2:-"32 64 128"}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
export B200DIST_PEER_TIMEOUT_S=60
for c in $CTAS; do
  timeout 170 $TR bench.py --gpus $N --steps 8 --warmup 3 --fsdp-ctas $c > gpurun_out/ctas${c}_n$N.out 2> gpurun_out/ctas${c}_n$N.err
  echo "ctas $c rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/ctas${c}_n$N.out | head -1) $(grep -o '"sm_mhz": [0-9]*' gpurun_out/ctas${c}_n$N.out)"
done
