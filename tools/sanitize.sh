#!/bin/bash
# compute-sanitizer over the kernels with hand-rolled synchronisation: the streamed sweeps (bulk-copy
# producer warp + mbarrier ring), the fused shared-memory coarse tail, the tiled blur, the BiCGStab loop and the
# allocation-free blur.  Run on the GPU box:   bash tools/sanitize.sh   -> gpurun_out/sanitize_{memcheck,racecheck}.log
# memcheck: out-of-bounds / misaligned accesses in global, shared and local memory.  racecheck: shared-memory hazards.
set -u
mkdir -p gpurun_out
SEL='streamed_smoother or multigrid_hierarchy or (test_solve_fp64_matches_direct_solution and mg) or rhs_and_blur or blur_noalloc_matches or true_residual_output'
for tool in memcheck racecheck; do
  timeout 900 /usr/local/cuda/bin/compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
      python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x -k "$SEL" > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool exit code $?" >> gpurun_out/sanitize_$tool.log
  tail -4 gpurun_out/sanitize_$tool.log
done
