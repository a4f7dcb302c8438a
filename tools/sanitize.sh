#!/bin/bash
# compute-sanitizer over the kernel tests (run on a GPU box; ONE tool per invocation -- on this
# driver several tools in one call have left the GPU unusable, see B200_PROFILING.md).
#   tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck]
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
# first make sure the same command passes without the tool
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "grad_dense or grad_ell or codegen" || exit 1
exec timeout 1800 compute-sanitizer --tool "$TOOL" --error-exitcode 3 \
  python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "grad_dense or grad_ell or codegen"
