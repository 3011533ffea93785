#!/bin/bash
# One GPU-box session: new parity tests, smoke, bench (+ reference arm), sanitizer, per-op trace.
# Everything lands in gpurun_out/. Usage: gpurun --timeout 2400 -- bash tools/gpu_round.sh
mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; nvidia-smi -L >> gpurun_out/host.txt
timeout 900 python -m pytest tests/test_7b_shapes_gpu.py tests/test_vs_reference_gpu.py tests/test_decode_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu --durations=8 > gpurun_out/t_new.log 2>&1; echo "rc=$?" >> gpurun_out/t_new.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "rc=$?" >> gpurun_out/bench_cfg2.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?" >> gpurun_out/bench_ref.err
timeout 300 python tools/step_trace.py 1 512 > gpurun_out/step_trace_512.txt 2>&1
timeout 600 compute-sanitizer --tool memcheck python -m pytest "tests/test_engine_gpu.py::test_persistent_step_kernel_vs_oracle_with_context[small-1-300]" -x -q -m gpu > gpurun_out/sanitizer_memcheck.log 2>&1; echo "rc=$?" >> gpurun_out/sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck python -m pytest "tests/test_engine_gpu.py::test_persistent_step_kernel_vs_oracle_with_context[small-1-300]" -x -q -m gpu > gpurun_out/sanitizer_racecheck.log 2>&1; echo "rc=$?" >> gpurun_out/sanitizer_racecheck.log
tail -3 gpurun_out/t_new.log gpurun_out/smoke.log gpurun_out/bench_cfg2.err gpurun_out/bench_ref.err gpurun_out/sanitizer_memcheck.log gpurun_out/sanitizer_racecheck.log
