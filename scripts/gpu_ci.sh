#!/bin/bash
# Runs on the GPU box under gpurun: staged so one faulting kernel cannot hide the other results.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/ci.log
  timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1
  echo "rc=$? $name" | tee -a gpurun_out/ci.log
  tail -n "${TAILN:-12}" "gpurun_out/$name.log"
}
STAGES=${STAGES:-"k_fp32 k_tc k_parity e2e_exact e2e_bf16 train smoke bench"}
# compute-sanitizer: ONE tool per gpurun call (B200_PROFILING.md), on a handful of small kernel cases that cover
# the hand-rolled mbarrier / TMEM / TMA pipelines:  STAGES=san_memcheck  /  STAGES=san_racecheck  /  ...
SAN_TESTS='tests/test_gpu_kernels.py -m gpu -q -x -k "(tcgen05_engine and (case0 or case1 or case12 or case18)) or (parity_engine and (case0 or case2 or case10)) or dual_prior or motion_kernels or split_bf16x3"'
SAN_TRAIN='tests/test_gpu_training.py -m gpu -q -x -k "(wgrad_kernel and (case0 or case8 or case11)) or flow_warp_backward"'
for s in $STAGES; do case $s in
  k_fp32) run k_fp32 150 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "fp32_engine or mixed or layout or se_layer or motion or dual_prior or round_and_bits or vector_path or frame_io" ;;
  k_probe) run k_probe 100 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "probe" ;;
  k_tc) run k_tc 150 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tcgen05" ;;
  k_parity) run k_parity 150 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "parity or split" ;;
  e2e_exact) run e2e_exact 400 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "exact_modes or batched" ;;
  train) run train 300 python -m pytest tests/test_gpu_training.py -m gpu -q ;;
  san_memcheck) run san_memcheck 900 bash -c "compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest $SAN_TESTS && compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest $SAN_TRAIN" ;;
  san_racecheck) run san_racecheck 900 bash -c "compute-sanitizer --tool racecheck --error-exitcode 7 python -m pytest $SAN_TESTS" ;;
  san_synccheck) run san_synccheck 900 bash -c "compute-sanitizer --tool synccheck --error-exitcode 7 python -m pytest $SAN_TESTS" ;;
  san_initcheck) run san_initcheck 900 bash -c "compute-sanitizer --tool initcheck --error-exitcode 7 python -m pytest $SAN_TESTS" ;;
  e2e_bf16) run e2e_bf16 300 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "bf16 or 1080p or pipeline" ;;
  smoke) run smoke 200 python -c "import __graft_entry__ as g; g.smoke()" ;;
  bench) run bench 400 python bench.py --steps 1 --warmup 2 --gop 8 ;;
  bench_full) run bench_full 1200 python bench.py ;;
  all_gpu) run all_gpu 1500 python -m pytest tests -m gpu -x -q ;;
esac; done
