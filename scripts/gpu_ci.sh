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
STAGES=${STAGES:-"k_fp32 k_tc e2e_fp32 e2e_bf16 smoke bench"}
for s in $STAGES; do case $s in
  k_fp32) run k_fp32 150 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "fp32_engine or mixed or layout or se_layer or motion or dual_prior or round_and_bits or vector_path or frame_io" ;;
  k_probe) run k_probe 100 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "probe" ;;
  k_tc) run k_tc 150 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tcgen05" ;;
  e2e_fp32) run e2e_fp32 300 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "fp32 or batched" ;;
  e2e_bf16) run e2e_bf16 200 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "bf16 or 1080p or pipeline" ;;
  smoke) run smoke 200 python -c "import __graft_entry__ as g; g.smoke()" ;;
  bench) run bench 400 python bench.py --steps 1 --warmup 2 --gop 8 ;;
  bench_full) run bench_full 1200 python bench.py ;;
  all_gpu) run all_gpu 1500 python -m pytest tests -m gpu -x -q ;;
esac; done
