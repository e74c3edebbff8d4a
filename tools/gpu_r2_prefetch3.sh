set -u
mkdir -p gpurun_out
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-40s %8.2f us/step kernel %.2f us frac %.3f' % (sys.argv[1], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5 --steps 30"
for rep in 1 2; do
for e in 65536 262144 1048576 4194304; do
for pn in 1 2; do
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs $e > gpurun_out/g.json 2>gpurun_out/g.err; one "h envs=$e next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs $e --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c envs=$e next=$pn" gpurun_out/g.json
done
done
timeout 300 $R1 --envs 1048576 --layout soa > gpurun_out/g.json 2>gpurun_out/g.err; one "h soa envs=1M" gpurun_out/g.json
timeout 300 $R1 --envs 1048576 --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa envs=1M" gpurun_out/g.json
timeout 300 $R1 --envs 262144 --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa envs=262144" gpurun_out/g.json
timeout 300 $R1 --envs 1048576 --robot unitree_g1 > gpurun_out/g.json 2>gpurun_out/g.err; one "g1 DPEnv walk envs=1M" gpurun_out/g.json
done
tail -2 gpurun_out/g.err
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
