set -u
mkdir -p gpurun_out
N=${NG:-2}
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-28s %8.2f us/step  total %7.1f us' % (sys.argv[1], d['ms_per_step']*1e3, d['ms_per_step']*1e3*d['steps']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
for k in 20 20 50 100 200; do
  timeout 200 python bench.py --no-cpu --no-e2e --no-sweep --steps $k --warmup 5 > gpurun_out/f1.json 2>/dev/null; one "N=1 steps=$k" gpurun_out/f1.json
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu --no-e2e --no-sweep --steps $k --warmup 5 > gpurun_out/fN.json 2>/dev/null; one "N=$N steps=$k" gpurun_out/fN.json
done
