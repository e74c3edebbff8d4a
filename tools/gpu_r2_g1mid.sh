set -u
mkdir -p gpurun_out
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 3 --steps 20 --robot unitree_g1 --combined"
run() { timeout 300 "$@" > gpurun_out/m.json 2> gpurun_out/m.err; echo "rc=$? $(python -c "import json;d=json.loads(open('gpurun_out/m.json').read().strip().splitlines()[-1]);print('%.1f us frac %.3f'%(d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))" 2>/dev/null)"; }
for e in 262144 1048576; do
echo -n "g1c $e default: "; run $R1 --envs $e
echo -n "g1c $e wide + 3 slots: "; DM_WIDE=2 DM_NBUF=3 run $R1 --envs $e
echo -n "g1c $e 3 capped groups: "; DM_WIDE=0 run $R1 --envs $e
done
