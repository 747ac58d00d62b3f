#!/bin/bash
# Runs the GPU test-suite in separate processes (a trap in one kernel must not hide the others).
# Usage (on the GPU box, from the repo root): bash tools/gpu_run_tests.sh [quick]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, pytest args...
  local name=$1; shift
  timeout -s KILL 900 python -m pytest "$@" -m gpu -q --tb=short -rA -p no:cacheprovider > gpurun_out/$name.log 2>&1
  echo "== $name: exit $?" | tee -a gpurun_out/summary.txt
  grep -E "^(FAILED|ERROR)" gpurun_out/$name.log | head -20 | tee -a gpurun_out/summary.txt
  tail -n 1 gpurun_out/$name.log | tee -a gpurun_out/summary.txt
}
: > gpurun_out/summary.txt
if [ "$1" != "quick" ]; then
  run audio tests/test_gpu_audio.py
  run kernels_small tests/test_gpu_kernels.py -k "not modconv3x3 and not up_convolution and not dual"
  run conv_fp32 tests/test_gpu_kernels.py -k "fp32 and (modconv3x3 or up_convolution)"
fi
run conv_tc tests/test_gpu_kernels.py -k "tcgen05 or dual or ((bf16 or fp16) and up_convolution)"
run synthesis tests/test_gpu_synthesis.py
run resize tests/test_gpu_resize.py
run pipeline tests/test_gpu_pipeline.py
run guards tests/test_gpu_guard_bands.py
