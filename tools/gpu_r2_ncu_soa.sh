set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep --envs 1048576 --layout soa"
$CMD > gpurun_out/plain_soa.log 2>&1 || { tail -20 gpurun_out/plain_soa.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_soa -s 4 -c 1 -o gpurun_out/soa_1M $CMD > gpurun_out/ncu_soa.log 2>&1
ls -la gpurun_out/*.ncu-rep
