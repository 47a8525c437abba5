#!/bin/bash
# Runs on the GPU box under gpurun: staged so one faulting kernel cannot hide the other results.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/ci.log
  timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1
  echo "rc=$? $name" | tee -a gpurun_out/ci.log
  tail -n 15 "gpurun_out/$name.log"
}
run k_fp32 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "fp32_engine or mixed or layout or se_layer or motion or dual_prior or round_and_bits"
run k_probe 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "probe"
run k_tc 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "tcgen05"
run e2e_fp32 900 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "fp32 or batched"
run e2e_bf16 600 python -m pytest tests/test_gpu_e2e.py -m gpu -q -s -k "bf16"
run smoke 600 python -c "import __graft_entry__ as g; g.smoke()"
run bench_bf16 900 python bench.py --steps 1 --warmup 1 --gop 8
