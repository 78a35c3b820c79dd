#!/bin/bash
# compute-sanitizer target for the single-GPU kernel tests:  tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck]
# Run on a GPU box from the repository root (one tool per invocation).  The multi-rank peer-memory kernels are covered
# by tests/dist_fused_check.py on a single rank (all "peers" local), which the sanitizer can follow.
set -e
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
python -m consensusml_b200._build > /dev/null
export CML_SANITIZE=1
exec compute-sanitizer --tool "$TOOL" --error-exitcode 9 --print-limit 20 \
  python -m pytest tests/test_lasso_gpu.py tests/test_gemm_gpu.py tests/test_normalize.py tests/test_de.py tests/test_trees.py \
                   tests/test_svm.py tests/test_stats.py tests/test_x_fused.py -x -q -m gpu
