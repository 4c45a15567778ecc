#!/bin/bash
# compute-sanitizer passes over the kernel unit tests.  One tool per gpurun call, smallest cases only:
#   gpurun --timeout 900 -- 'tools/sanitize.sh memcheck'      (also: racecheck | synccheck | initcheck)
# The unit tests use batches of 37-200 images, which keeps a memcheck pass in the minutes range; the large-batch cases
# (-k "not 777 and not 640 and not 600") are excluded.  Output: gpurun_out/sanitize_<tool>.log.
# Not run in round 1: the GPU budget went to measurements; the peer-memory barrier protocol (flag release/acquire) and the
# mbarrier pipelines are the surfaces to look at first.
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 850 compute-sanitizer --tool "$TOOL" --error-exitcode 7 --launch-timeout 0 \
  python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "not 777 and not 640 and not 600" \
  > "gpurun_out/sanitize_${TOOL}.log" 2>&1
rc=$?
tail -15 "gpurun_out/sanitize_${TOOL}.log"
exit $rc
