#!/bin/bash
# One GPU session: parity tests, bench, ncu launch list, one full ncu capture of the step kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench_exit=$?" >> gpurun_out/bench.err
SMALL="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-sweep"
$SMALL > gpurun_out/plain_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_small.log 2>&1
BIG="python bench.py --envs 1048576 --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep ${BIG_EXTRA:-}"
$BIG > gpurun_out/plain_big.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:step_ -s 3 -c 2 -f -o gpurun_out/prof_step $BIG > gpurun_out/ncu_big.log 2>&1
echo done
