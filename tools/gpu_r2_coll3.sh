set -u
mkdir -p gpurun_out
N=${NG:-2}
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d['config'].get('collective')
    print('%-34s n=%d  %8.2f us/step  value %.3e  kernel %.2f us  coll=%s count=%s/%s' % (sys.argv[1], d['n_gpus'], d['ms_per_step']*1e3, d['value'], d['roofline']['kernel_ms']*1e3,
          (c or {}).get('mode'), d.get('stats',{}).get('count'), d.get('stats',{}).get('count_expected')))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[2].replace('.json','.err')).read()[-1500:])
PY
}
echo "== multirank test"; timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu 2>&1 | tail -5
Q="--no-cpu --no-e2e --no-sweep --steps 200 --warmup 20"
timeout 200 python bench.py $Q > gpurun_out/e1.json 2> gpurun_out/e1.err; show "N=1 (no collective)" gpurun_out/e1.json
for envs in 4096 65536 1048576; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N $Q --envs $envs > gpurun_out/eN_$envs.json 2> gpurun_out/eN_$envs.err; show "N=$N auto envs=$envs" gpurun_out/eN_$envs.json
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/eN_drv.json 2> gpurun_out/eN_drv.err; show "N=$N driver-style (20 steps, full)" gpurun_out/eN_drv.json
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/eN_drv.json').read().strip().splitlines()[-1])
    print('e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
    for k,v in d['configs'].items():
        if k!='config5_sweep': print(' ', k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('envs_per_gpu','ms_per_step','kernel_ms','frac','env_steps_per_s','collective')})
except Exception as e: print('ERR', e)
PY
