#!/bin/bash
# The cross-GPU tests (one-shot NVLink allreduce protocol, GEMV-epilogue push, TP decode parity) need >= 2 GPUs and
# skip on a one-GPU box. Run on a multi-GPU lease:  bash tests/run_multi_gpu.sh [log]
# (each test spawns its own ranks with torch.multiprocessing over 127.0.0.1; no torchrun needed)
LOG=${1:-gpurun_out/multi_gpu_tests.log}
mkdir -p "$(dirname "$LOG")"
nvidia-smi -L | tee "$LOG"
python -m pytest tests/test_tp_gpu.py -q -m gpu -rs 2>&1 | tee -a "$LOG"
exit ${PIPESTATUS[0]}
