# final form of the collective: multirank parity (5 routes), then the bench line at N = 1 and N = 2 (driver's command line)
set -u
mkdir -p gpurun_out
N=${NG:-2}
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d.get('stats',{}).get('count')
    print('%-44s %8.2f us/step  total %7.1f us  count %s/%s launches %s kernel_ms %.4f' % (sys.argv[1], d['ms_per_step']*1e3, d['ms_per_step']*1e3*d['steps'], c, d['n_gpus']*4096*d['steps'], d.get('gpu_launches'), d['roofline']['kernel_ms']))
    for k,v in (d.get('configs') or {}).items():
        if isinstance(v,dict): print('     ',k,{a:b for a,b in v.items() if a in('envs_per_gpu','ms_per_step','kernel_ms','frac')}, (v.get('collective') or {}).get('count'))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q 2>&1 | tail -3
R1="python bench.py --gpus 1 --no-cpu --no-e2e --warmup 5"
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu --no-e2e --warmup 5"
for rep in 1 2; do
timeout 300 $R1 --steps 20 --no-configs > gpurun_out/g.json 2>gpurun_out/g.err; one "N=1 steps=20" gpurun_out/g.json
timeout 300 $R --steps 20 --no-configs > gpurun_out/g.json 2>gpurun_out/g.err; one "N=$N steps=20" gpurun_out/g.json
timeout 300 $R --steps 200 --no-configs > gpurun_out/g.json 2>gpurun_out/g.err; one "N=$N steps=200" gpurun_out/g.json
done
timeout 600 $R --steps 20 > gpurun_out/g2full.json 2>gpurun_out/g2full.err; one "N=$N steps=20 with configs" gpurun_out/g2full.json
tail -3 gpurun_out/g2full.err
