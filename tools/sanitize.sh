#!/usr/bin/env bash
# Race / memory checking of the native kernels with compute-sanitizer (the reference ships no sanitizer
# recipe).  Single GPU only: the sanitizer serialises kernels, so multi-rank flag protocols would hit their watchdog.
#   tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
# memcheck : out-of-bounds / misaligned global, shared and TMEM-adjacent accesses
# racecheck: shared-memory hazards between the warps of a CTA (producer / MMA / epilogue hand-offs)
# synccheck: illegal barrier usage (mbarrier, __syncthreads under divergence)
set -euo pipefail
TOOL=${1:-memcheck}; shift || true
EXPR=${1:-"norms or elementwise or simt"}   # test ids are the group names of tests/test_gpu_kernels.py
cd "$(dirname "$0")/.."
export B200DIST_DEBUG_SYNC=1
exec compute-sanitizer --tool "$TOOL" --error-exitcode 9 --kernel-name-exclude kns=at:: \
     python -m pytest tests/test_gpu_kernels.py -q -x -m gpu -k "$EXPR"
