#!/bin/bash
# Quick GPU check: parity tests, then the bench without the CPU / e2e legs (kernel numbers + sweep).
set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --no-cpu --no-e2e ${BENCH_EXTRA:-} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench_exit=$?"
cat gpurun_out/bench_quick.json
if [ "${SANITIZE:-0}" = "1" ]; then
  timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_batch.py -x -q -m gpu -k "tile_kernel" > gpurun_out/sanitize.log 2>&1
  tail -5 gpurun_out/sanitize.log
fi
