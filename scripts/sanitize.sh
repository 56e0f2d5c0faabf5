#!/bin/bash
# compute-sanitizer over the small-shape kernel tests (one tool per invocation, see
# B200_PROFILING.md).  usage: scripts/sanitize.sh [memcheck|racecheck|initcheck|synccheck]
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$TOOL" --error-exitcode 9 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x \
     -k "shift_exact or pagerank_100 or heat_512 or sanitize or vigenere or dgemm or u8_to_bf16"
