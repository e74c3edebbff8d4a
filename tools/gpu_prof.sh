#!/bin/bash
# usage: gpu_prof.sh <tag> <bench args...>   -> gpurun_out/prof_<tag>.ncu-rep (full set, step kernel only)
set -u
tag=$1; shift
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep $*"
$CMD > gpurun_out/plain_$tag.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:step_ -s 4 -c 1 -f -o gpurun_out/prof_$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
echo "prof $tag done"
