#!/bin/bash
# One GPU session: parity tests, bench, ncu launch list, full ncu captures of the step kernel reduced to text
# summaries ON THE BOX (the .ncu-rep files are too big to bring back: gpurun_out/ is capped at 64 MiB).
set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench_exit=$?" >> gpurun_out/bench.err
SMALL="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-sweep"
$SMALL > gpurun_out/plain_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_small.log 2>&1
prof() {  # prof <tag> <bench args...>: plain run first, then one full-set capture of the step kernel, summarised
  local tag=$1; shift
  local CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep $*"
  $CMD > gpurun_out/plain_$tag.log 2>&1 || return
  ncu --set full --clock-control none --import-source on -k regex:step_ -s 4 -c 1 -f -o /tmp/prof_$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
  { echo "# $CMD"; echo "# ncu --set full --clock-control none --import-source on -k regex:step_ -s 4 -c 1"; python tools/ncu_summary.py /tmp/prof_$tag.ncu-rep;
    echo; echo "# per-instruction stall samples (SASS view)"; python tools/ncu_hot.py /tmp/prof_$tag.ncu-rep | head -40; } > gpurun_out/summary_$tag.txt 2>&1
  rm -f /tmp/prof_$tag.ncu-rep
}
prof h1M --envs 1048576
prof h4096 --envs 4096
prof g1c_1M --robot unitree_g1 --combined --envs 1048576
prof g1c_262144 --robot unitree_g1 --combined --envs 262144
echo done
