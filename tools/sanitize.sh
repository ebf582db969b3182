#!/bin/bash
# compute-sanitizer over a reduced -m gpu subset that covers the hand-rolled synchronisation: the cooperative / cluster
# dopri5 kernels (global spin barrier, DSMEM slot exchange), certify_tc_kernel (mbarrier ring, setmaxnreg), gemm_tc_kernel,
# the Gauss-Jordan kernels and the fused backbone.  Logs under gpurun_out/<tag>/; summaries are copied to profiles/ by hand.
#   gpurun --timeout 1500 -- 'bash tools/sanitize.sh r02_sanitizer'
TAG=${1:-sanitizer}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SUBSET='batched_cluster_solver_equals_single_solves and (3-64 or 2-96) or certify_tensor_core_path or linear_tensor_core_path and (128-64 or 300-512) or test_cayley_linear and not bwd or backbone_fused or ivp_rows_per_cta_invariance or projection_forward_backward or segway_simulate or fi_loss_forward_backward and uniform'
for TOOL in memcheck racecheck synccheck; do
  timeout 420 compute-sanitizer --tool $TOOL --log-file $OUT/$TOOL.log --print-limit 20 \
    python -m pytest tests -m gpu -x -q -k "$SUBSET" -p no:cacheprovider > $OUT/$TOOL.pytest.log 2>&1
  echo "$TOOL rc=$?" | tee -a $OUT/rc.txt
  tail -3 $OUT/$TOOL.pytest.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" $OUT/$TOOL.log | tail -5
done
