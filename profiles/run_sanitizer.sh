#!/bin/bash
# compute-sanitizer over the warp-specialised kernels' unit tests (SURVEY 5 build note).
#   gpurun --timeout 2400 -- 'bash profiles/run_sanitizer.sh'   ->  gpurun_out/r02_sanitizer_{memcheck,racecheck}.txt
# Small batches only: every kernel runs 10-100x slower under the tools.
set -u
mkdir -p gpurun_out
SEL='test_deepfm_whole_kernel and (100- or 1000-) or test_dcn_whole_kernel and (900 or 130 or 129) or test_shapes or test_widedeep_one_launch and 129'
for tool in memcheck racecheck; do
  out=gpurun_out/r02_sanitizer_$tool.txt
  echo "== compute-sanitizer --tool $tool, pytest -k \"$SEL\"" > $out
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 77 \
      python -m pytest tests/test_gpu_fused.py tests/test_gpu_din_tc.py -m gpu -q -x -k "$SEL" >> $out 2>&1
  echo "== exit code $? (77 = the tool reported errors, 124 = timeout)" >> $out
  tail -15 $out
done
