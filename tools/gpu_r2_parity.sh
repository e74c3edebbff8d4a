set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_baseline_configs.py -x -q -m gpu --durations=5 2>&1 | tail -25
