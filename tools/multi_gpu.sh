#!/bin/bash
# usage: multi_gpu.sh N   -> gpurun_out/bench_${N}gpu_4096.json, bench_${N}gpu_1M_allreduce_every_step.json
N=$1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$T bench.py --gpus $N --steps 200 --warmup 20 --no-sweep 2> gpurun_out/bench_${N}gpu_4096.err | tail -1 > gpurun_out/bench_${N}gpu_4096.json
$T bench.py --gpus $N --steps 100 --warmup 10 --envs 1048576 --allreduce-every 1 --no-sweep --no-e2e 2> gpurun_out/bench_${N}gpu_1M.err | tail -1 > gpurun_out/bench_${N}gpu_1M_allreduce_every_step.json
python - <<PY
import json
for f in ("gpurun_out/bench_${N}gpu_4096.json", "gpurun_out/bench_${N}gpu_1M_allreduce_every_step.json"):
    try:
        d = json.loads(open(f).read())
        e = d.get("e2e") or {}
        print(f, "value %.3e ms/step %.5f n_gpus %d" % (d["value"], d["ms_per_step"], d["n_gpus"]), "e2e %.3e" % e.get("value", 0), (e.get("pinned") or {}).get("value"))
    except Exception as ex:
        print(f, "FAILED", ex)
PY
