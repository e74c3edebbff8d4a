set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
DM_HOST_TRACE=1 python tools/e2e_scale_probe.py 262144 1 > gpurun_out/e2e_scale_g1c.log 2>&1; cat gpurun_out/e2e_scale_g1c.log
DM_HOST_TRACE=1 python tools/e2e_scale_probe.py 65536 0 > gpurun_out/e2e_scale_h64k.log 2>&1; cat gpurun_out/e2e_scale_h64k.log
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; lscpu | head -30 > gpurun_out/lscpu.txt; nvidia-smi -q | grep -i -A6 "pci\b\|Link" | head -60 > gpurun_out/pcie.txt
