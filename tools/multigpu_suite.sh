#!/usr/bin/env bash
# Measurement suite for N GPUs of one box (both arms of the headline benchmarks, collectives, timeline).
#   tools/multigpu_suite.sh N [llama|resnet|comm|hybrid|pp ...]     results -> gpurun_out/*_nN.*
# Every step runs under its own `timeout`; a hung step cannot eat the whole lease.
set -u
N=${1:-8}; shift || true
WHAT=${*:-"llama resnet comm"}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export B200DIST_PEER_TIMEOUT_S=${B200DIST_PEER_TIMEOUT_S:-60}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run() { local name=$1 secs=$2; shift 2; timeout "$secs" "$@" > "gpurun_out/${name}_n${N}.out" 2> "gpurun_out/${name}_n${N}.err"; echo "== $name rc=$?"; tail -c 900 "gpurun_out/${name}_n${N}.out"; echo; }
for w in $WHAT; do
  case $w in
    llama)
      run bench 170 $TR bench.py --gpus $N --steps 10 --warmup 3
      run bench_ref 230 $TR bench.py --impl reference --gpus $N --steps 10 --warmup 3
      run timeline 130 $TR tools/fsdp_timeline.py ;;
    resnet)
      run resnet 130 $TR bench/resnet_bench.py --gpus $N --steps 10 --warmup 3
      run resnet_ref 160 $TR bench/resnet_bench.py --gpus $N --impl reference --steps 10 --warmup 3 ;;
    comm)
      run comm_check 170 $TR tools/comm_check.py
      run comm_bench 260 $TR bench/comm_bench.py --impl both --benchmark_runs 10 --warmup_runs 3 --output_file gpurun_out/comm_bench_${N}gpu.csv ;;
    hybrid)
      run hybrid_fused 200 $TR bench/hybrid_bench.py --tp 4 --fused --steps 6 --warmup 3
      run hybrid 200 $TR bench/hybrid_bench.py --tp 4 --no-fused --steps 6 --warmup 3
      run hybrid_ref 260 $TR bench/hybrid_bench.py --tp 4 --impl reference --steps 6 --warmup 3 ;;
    pp)
      run pp 200 $TR bench/pp_bench.py --pp 4 --dp 2
      run pp_ref 240 $TR bench/pp_bench.py --pp 4 --dp 2 --impl reference ;;
    parallel)
      run parallel_check 400 $TR tools/parallel_check.py ;;
  esac
done
