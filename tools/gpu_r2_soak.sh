# soak: many consecutive steps from one seeded state at sizes where every SM carries many tiles
set -u
mkdir -p gpurun_out
R="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5"
run() { timeout 600 $R "$@" > gpurun_out/soak.json 2> gpurun_out/soak.err; echo "rc=$? $* $(grep -o 'line [0-9]*, in [a-z_]*' gpurun_out/soak.err | tail -1) $(python -c "import json;d=json.loads(open('gpurun_out/soak.json').read().strip().splitlines()[-1]);print('%.1f us'%(d['ms_per_step']*1e3), d['stats']['count']==d['stats']['count_expected'])" 2>/dev/null)"; }
run --steps 400 --envs 1048576
run --steps 400 --envs 1048576 --layout soa
run --steps 200 --envs 1048576 --robot unitree_g1 --combined
run --steps 200 --envs 1048576 --robot unitree_g1 --combined --layout soa
run --steps 100 --envs 4194304
run --steps 1000 --envs 262144
run --steps 1000 --envs 65536 --motion spinkick,dance_b
run --steps 3000 --envs 4096
