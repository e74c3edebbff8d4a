# flag form vs tagged form of the statistics exchange (2 GPUs)
set -u
mkdir -p gpurun_out
N=${NG:-2}
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d.get('stats',{}).get('count')
    print('%-44s %8.2f us/step  total %7.1f us  count %s/%s' % (sys.argv[1], d['ms_per_step']*1e3, d['ms_per_step']*1e3*d['steps'], c, d['n_gpus']*4096*d['steps']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu --no-e2e --no-sweep --no-configs --warmup 5"
DM_PEER_PULL=2 timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q 2>&1 | tail -3
for rep in 1 2; do
for st in 20 200; do
DM_BENCH_GATE_US=300 timeout 300 $R --steps $st > gpurun_out/g.json 2>/dev/null; one "N=$N steps=$st flags gate" gpurun_out/g.json
DM_BENCH_GATE_US=300 DM_PEER_PULL=2 timeout 300 $R --steps $st > gpurun_out/g.json 2>/dev/null; one "N=$N steps=$st tagged gate" gpurun_out/g.json
DM_PEER_PULL=2 timeout 300 $R --steps $st > gpurun_out/g.json 2>/dev/null; one "N=$N steps=$st tagged" gpurun_out/g.json
DM_PEER_DEBUG=3 DM_BENCH_GATE_US=300 timeout 300 $R --steps $st > gpurun_out/g.json 2>/dev/null; one "N=$N steps=$st extra CTA only waits, gate" gpurun_out/g.json
done
done
nvidia-smi topo -m 2>&1 | head -20
lscpu | grep -i "numa\|socket\|model name\|^CPU(s)" 
