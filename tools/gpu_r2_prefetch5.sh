set -u
mkdir -p gpurun_out
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-44s %8.2f us/step kernel %.2f us frac %.3f' % (sys.argv[1], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5"
for rep in 1 2; do
for pn in 0 1 2; do
for st in 10 30; do
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps $st --envs 1048576 --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa 1M next=$pn steps=$st" gpurun_out/g.json
done
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps 20 --envs 262144 --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa 262144 next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps 20 --envs 1048576 --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c aos 1M next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps 20 --envs 262144 --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c aos 262144 next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps 20 --envs 65536 --motion spinkick,dance_b > gpurun_out/g.json 2>gpurun_out/g.err; one "c3 aos 65536 next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --steps 20 --envs 1048576 --layout soa --robot unitree_g1 > gpurun_out/g.json 2>gpurun_out/g.err; one "g1 DPEnv soa 1M next=$pn" gpurun_out/g.json
done
done
tail -2 gpurun_out/g.err
