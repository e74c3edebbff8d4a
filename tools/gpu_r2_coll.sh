set -u
mkdir -p gpurun_out
N=${NG:-2}
echo "== multirank test"; timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu 2>&1 | tail -15
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d['config'].get('collective')
    print('%-34s n=%d  %8.2f us/step  value %.3e  kernel %.2f us  coll=%s count=%s/%s' % (sys.argv[1], d['n_gpus'], d['ms_per_step']*1e3, d['value'], d['roofline']['kernel_ms']*1e3,
          (c or {}).get('mode'), d.get('stats',{}).get('count'), d.get('stats',{}).get('count_expected')))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[2].replace('.json','.err')).read()[-1800:])
PY
}
Q="--no-cpu --no-e2e --no-sweep --steps 200 --warmup 20"
for m in off auto reduce; do
  timeout 200 python bench.py $Q --collective $m --collective-at-1 > gpurun_out/c1_$m.json 2> gpurun_out/c1_$m.err; show "N=1 collective=$m" gpurun_out/c1_$m.json
done
for m in auto reduce nccl; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N $Q --collective $m > gpurun_out/cN_$m.json 2> gpurun_out/cN_$m.err; show "N=$N collective=$m" gpurun_out/cN_$m.json
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N $Q --collective auto --envs 1048576 > gpurun_out/cN_1M.json 2> gpurun_out/cN_1M.err; show "N=$N 1M auto" gpurun_out/cN_1M.json
echo "== default bench N=1 (configs block)"; T0=$(date +%s); timeout 900 python bench.py --no-cpu > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench wall $(( $(date +%s) - T0 )) s"; python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
    print('value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
    for k,v in d['configs'].items():
        if k=='config5_sweep':
            for r in v: print('  sweep', r['envs_per_gpu'], round(r['humanoid3d_walk']['kernel_ms']*1e3,2), round(r['humanoid3d_walk']['frac'],3), round(r['g1_combined']['kernel_ms']*1e3,2), round(r['g1_combined']['frac'],3))
        else: print(' ', k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms_per_step','kernel_ms','frac','env_steps_per_s','e2e')})
except Exception as e:
    print('ERR', e); print(open('gpurun_out/bench_full.err').read()[-2500:])
PY
