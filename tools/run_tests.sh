#!/usr/bin/env bash
# Communication smoke tests on one box - the role of the reference's tests/run_tests.sh and tests/pbs_run_tests.sh (which
# submit 2-node PBS jobs and repeat each test with nccl / gloo and with the aws-ofi-nccl plugin on and off; neither PBS nor
# the plugin exists on a single NVSwitch box, so the matrix here is backend x {p2p, collectives, environment}).
#   tools/run_tests.sh [nproc]      (default: all visible GPUs, or 2 CPU processes)
set -euo pipefail
cd "$(dirname "$0")/.."
NGPU=$(python -c "import torch; print(torch.cuda.device_count())")
NPROC=${1:-$([ "$NGPU" -gt 0 ] && echo "$NGPU" || echo 2)}
BACKENDS=$([ "$NGPU" -gt 0 ] && echo "nccl gloo" || echo "gloo")
python tools/print_hostinfo.py
python tools/test_env.py
for be in $BACKENDS; do
  echo "=== backend $be, $NPROC processes ==="
  export B200DIST_BACKEND=$be
  scripts/launch.sh "$NPROC" tools/test_torchrun.py
  scripts/launch.sh "$NPROC" tools/send_recv_test.py
  scripts/launch.sh "$NPROC" tools/all_reduce_test.py --backend "$be" --max-exp 6 --output results/benchmark_results.log
done
scripts/launch.sh "$NPROC" tools/check_environment.py
if [ "$NGPU" -ge 2 ]; then
  scripts/launch.sh "$NPROC" tools/comm_check.py
  scripts/launch.sh "$NPROC" bench/comm_bench.py --impl both --output_file results/benchmark_results.log
fi
