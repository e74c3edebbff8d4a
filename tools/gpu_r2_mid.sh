# the middle shape (3 groups, 4 ring slots) after the slot-tenant fix: soak at sizes that failed before, then timings
set -u
mkdir -p gpurun_out
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 3"
run() { timeout 600 "$@" > gpurun_out/m.json 2> gpurun_out/m.err; echo "rc=$? $(grep -m1 -o 'CUDA error: [a-z ]*\|AssertionError.*' gpurun_out/m.err) $(python -c "import json;d=json.loads(open('gpurun_out/m.json').read().strip().splitlines()[-1]);print('%.1f us frac %.3f count ok %s'%(d['roofline']['kernel_ms']*1e3, d['roofline']['frac'], d['stats']['count']==d['stats']['count_expected']))" 2>/dev/null)"; }
echo -n "h 4M steps=60: "; run $R1 --steps 60 --envs 4194304
echo -n "h 3M steps=60: "; run $R1 --steps 60 --envs 3145728
echo -n "h 2M steps=300: "; run $R1 --steps 300 --envs 2097152
echo -n "h 1M steps=400: "; run $R1 --steps 400 --envs 1048576
echo -n "c3 4M steps=40: "; run $R1 --steps 40 --envs 4194304 --motion spinkick,dance_b
echo -n "h 4M old bound (4 capped groups) steps=20: "; DM_MID_HI=80 run $R1 --steps 20 --envs 4194304
echo -n "h 262144 steps=200: "; run $R1 --steps 200 --envs 262144
echo -n "h 1M NBUF=6 steps=60: "; DM_NBUF=6 run $R1 --steps 60 --envs 1048576
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
