#!/bin/bash
# compute-sanitizer on small shapes (one tool per gpurun call, see /opt/skills/guides/B200_PROFILING.md):
#   gpurun -- 'bash scripts/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 3 python -m pytest -x -q -m gpu \
    tests/test_gpu_statevector.py::test_groups_and_autograd tests/test_gpu_train_agg.py::test_param_shift_training_equals_adjoint_training \
    "tests/test_gpu_train_agg.py::test_fused_aggregation_matches_reference[delta-0.7-True-4-96]" \
    > gpurun_out/sanitizer_$TOOL.log 2>&1
echo "sanitizer $TOOL rc=$?"; tail -15 gpurun_out/sanitizer_$TOOL.log
