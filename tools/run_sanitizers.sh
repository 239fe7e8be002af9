#!/bin/bash
# compute-sanitizer over the hand-written kernels.  ONE tool per invocation -- and per gpurun call:
#   tools/run_sanitizers.sh memcheck | racecheck | synccheck | initcheck
# The conv kernel aliases its epilogue staging, BatchNorm column scratch and finalize scratch onto drained pipeline
# stages and hand-rolls mbarrier / TMEM / cluster-barrier protocols: racecheck and synccheck are aimed at exactly that.
# The subset is small on purpose (the tools slow kernels down 10-100x): one case per code path -- unsplit, cluster
# split-K (DSMEM reduce), workspace split-K, folded BatchNorm finalize, fused head, BatchNorm backward, emulated-rank
# collectives.  Summary lines land in gpurun_out/sanitizer_<tool>.txt; copy them to profiles/ when they are clean.
set -u
TOOL=${1:-memcheck}
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
LOG=$OUT/sanitizer_$TOOL.log
SUM=$OUT/sanitizer_$TOOL.txt
TESTS="tests/test_gpu_conv.py tests/test_gpu_ops.py tests/test_gpu_comm.py"
# one case per code path: unsplit, cluster split-K (DSMEM reduce), workspace split-K, conv -> BatchNorm through the
# accumulators, fused head, BatchNorm backward (both paths), first layer on the tensor cores, emulated-rank collectives
SELECT="test_conv_tiles_and_split_k and (128-1 or 128-8 or 64-16 or 128-3) or test_conv_then_batchnorm_from_accumulators and 8-4-256-512 \
or test_fused_head_forward_backward or test_bn_relu_pool_forward_backward and 32-512-2 or test_bn_backward_separate_finalize_kernel and 8-128-16 \
or test_first_conv_fprop_and_wgrad and 8-32-1 or test_allreduce_oneshot_twoshot_and_gather and 4"
EXTRA=""
[ "$TOOL" = "racecheck" ] && EXTRA="--racecheck-report all"
echo "compute-sanitizer --tool $TOOL $EXTRA : $(date -u)" > "$SUM"
timeout 3000 compute-sanitizer --tool "$TOOL" $EXTRA --error-exitcode 99 --print-limit 30 \
  python -m pytest $TESTS -k "$SELECT" -x -q -p no:cacheprovider > "$LOG" 2>&1
RC=$?
echo "exit code: $RC" >> "$SUM"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" "$LOG" | tail -20 >> "$SUM"
grep -E "^=========     (at|in) |Race reported|Invalid|Barrier error|hazard" "$LOG" | sort | uniq -c | sort -rn | head -40 >> "$SUM"
cat "$SUM"
exit $RC
