#!/bin/bash
# The round's closing GPU session: gpu_round.sh (tests, bench, launch list, ncu summaries) + size sweep + the per-GPU
# shards of BASELINE configs 3 / 4 + the host-step probe.
bash tools/gpu_round.sh
python tools/sweep.py --envs 1024,2048,4096,8192,16384,32768,65536,131072,262144,524288,1048576,2097152,4194304 --cases h,c --stats 1 --steps 50 2>/dev/null | grep "^{" > gpurun_out/sweep_final.jsonl
python bench.py --motion spinkick,dance_b --envs 8192 --no-cpu --no-sweep --no-e2e 2>/dev/null | tail -1 > gpurun_out/bench_cfg3_shard.json
python bench.py --robot unitree_g1 --combined --envs 32768 --no-cpu --no-sweep --no-e2e 2>/dev/null | tail -1 > gpurun_out/bench_cfg4_shard.json
python bench.py --robot unitree_g1 --combined --envs 262144 --no-cpu --no-sweep 2>/dev/null | tail -1 > gpurun_out/bench_cfg4_1gpu.json
python tools/zerocopy_probe.py 4096 16384 65536 2>&1 | grep -v Warning > gpurun_out/zerocopy_final.log
