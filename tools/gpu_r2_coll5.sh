set -u
mkdir -p gpurun_out
N=${NG:-2}
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-40s %8.2f us/step  total %7.1f us' % (sys.argv[1], d['ms_per_step']*1e3, d['ms_per_step']*1e3*d['steps']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu --no-e2e --no-sweep --warmup 5"
for rep in 1 2; do
timeout 300 $R --steps 20 > gpurun_out/g.json 2>/dev/null; one "N=$N steps=20 default" gpurun_out/g.json
DM_BENCH_NO_FINAL=1 timeout 300 $R --steps 20 > gpurun_out/g.json 2>/dev/null; one "N=$N steps=20 NO_FINAL" gpurun_out/g.json
DM_BENCH_NO_ALIGN=1 timeout 300 $R --steps 20 > gpurun_out/g.json 2>/dev/null; one "N=$N steps=20 NO_ALIGN" gpurun_out/g.json
DM_BENCH_NO_FINAL=1 DM_BENCH_NO_ALIGN=1 timeout 300 $R --steps 20 > gpurun_out/g.json 2>/dev/null; one "N=$N steps=20 NO_FINAL NO_ALIGN" gpurun_out/g.json
timeout 300 $R --steps 20 --collective off > gpurun_out/g.json 2>/dev/null; one "N=$N steps=20 collective off" gpurun_out/g.json
done
