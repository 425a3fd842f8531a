#!/bin/sh
# compute-sanitizer over the hand-written mbarrier / setmaxnreg / TMA / tcgen05 pipelines, on small shapes:
#     sh tools/sanitize.sh            (on the GPU box; writes gpurun_out/sanitize_*.log and a summary)
# memcheck: out-of-bounds / misaligned accesses of every kernel the selected tests launch;
# racecheck: shared-memory hazards between the producer / consumer / epilogue warps (the full / empty / staged /
#            drained mbarrier hand-shakes of contract.cu, contract_f32.cu, contract_tc05.cu);
# synccheck: barrier misuse (named barriers of the producer warpgroups, __syncthreads under divergence).
# The selection keeps the run to a few minutes: the raw contraction cases (all kernel variants, both dtypes),
# the golden residuals and one (T) test.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SEL='test_contract_vs_einsum or test_contract_f32_tcgen05 or test_residuals_golden or test_golden_triples_on_test_amplitudes or test_dot_many'
SUMMARY=gpurun_out/sanitize_summary.txt
: > $SUMMARY
for tool in memcheck racecheck synccheck; do
  log=gpurun_out/sanitize_$tool.log
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
      python -m pytest tests/test_gpu_kernels.py tests/test_gpu_solver.py -m gpu -q -x -k "$SEL" > $log 2>&1
  rc=$?
  echo "== $tool: exit code $rc" >> $SUMMARY
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|hazard" $log | tail -6 >> $SUMMARY
done
cat $SUMMARY
