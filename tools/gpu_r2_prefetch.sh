set -u
mkdir -p gpurun_out
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-34s %8.2f us/step kernel %.2f us' % (sys.argv[1], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5"
for rep in 1 2; do
for e in 4096 16384 65536; do
timeout 300 $R1 --steps 50 --envs $e > gpurun_out/g.json 2>gpurun_out/g.err; one "h envs=$e prefetch" gpurun_out/g.json
DM_NO_PREFETCH=1 timeout 300 $R1 --steps 50 --envs $e > gpurun_out/g.json 2>gpurun_out/g.err; one "h envs=$e no prefetch" gpurun_out/g.json
done
timeout 300 $R1 --steps 50 --envs 8192 --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c envs=8192 prefetch" gpurun_out/g.json
DM_NO_PREFETCH=1 timeout 300 $R1 --steps 50 --envs 8192 --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c envs=8192 no prefetch" gpurun_out/g.json
done
tail -2 gpurun_out/g.err
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
