#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests.
#
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
#
# NOTE: the pool this repo was developed on refuses compute-sanitizer ("closed on this pool"), so the checks
# that run everywhere are the repo's own: HVD_DEBUG_GUARD=1 red zones (MnistEngine.check_guards), HVD_DEBUG_POISON=1
# NaN-poisoned gradients, HVD_DEBUG_CHECKSUM=N replica checksums, and the CPU-oracle comparisons in tests/.
# This script is the recipe for boxes where the tool is allowed.
# One tool per invocation (and per gpurun call).  The caching allocator is turned off so an out-of-bounds
# access inside a torch pool block is not hidden; the barrier-timeout test is left out because it traps on
# purpose.  The report lands in gpurun_out/sanitizer_<tool>.txt; copy the summary to profiles/.
set -u
tool=${1:-memcheck}
sel=${2:-"tc_gemm or conv1_fwd or conv2 or head or dense1 or dense_forward or emulated"}
out=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$out"
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
export HVD_NO_GRAPH=1
timeout "${SANITIZE_TIMEOUT:-600}" compute-sanitizer --tool "$tool" --error-exitcode 77 --print-limit 20 \
  --log-file "$out/sanitizer_${tool}.txt" \
  python -m pytest tests/test_gpu_kernels.py tests/test_gpu_collectives.py tests/test_zz_no_alloc.py \
  -x -q -p no:cacheprovider -k "($sel) and not timeout and not multi_gpu" 2>&1 | tail -5
rc=${PIPESTATUS[0]}
echo "sanitizer($tool) rc=$rc"
tail -5 "$out/sanitizer_${tool}.txt"
exit 0
