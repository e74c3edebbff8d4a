set -u
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d['config'].get('collective') or {}
    e=d.get('e2e',{})
    print('%-22s n=%d  %8.2f us/step  value %.3e  coll=%s count=%s/%s  e2e %.3e (%s cpus/rank) pinned %.3e' % (sys.argv[1], d['n_gpus'], d['ms_per_step']*1e3, d['value'],
          c.get('mode'), d.get('stats',{}).get('count'), d.get('stats',{}).get('count_expected'), e.get('value',0), e.get('cpus_per_rank'), e.get('pinned',{}).get('value',0)))
    for k,v in d.get('configs',{}).items():
        if k!='config5_sweep': print('    ', k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('envs_per_gpu','ms_per_step','kernel_ms','frac','env_steps_per_s')})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[2].replace('.json','.err')).read()[-1500:])
PY
}
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-configs > gpurun_out/s1.json 2> gpurun_out/s1.err; show "N=1" gpurun_out/s1.json
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/s$N.json 2> gpurun_out/s$N.err; show "N=$N" gpurun_out/s$N.json
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 200 --warmup 20 --no-e2e --no-sweep > gpurun_out/s8_200.json 2> gpurun_out/s8_200.err; show "N=8 200 steps" gpurun_out/s8_200.json
cp gpurun_out/s8.json gpurun_out/r02_bench_8gpu.json; cp gpurun_out/s4.json gpurun_out/r02_bench_4gpu.json; cp gpurun_out/s2.json gpurun_out/r02_bench_2gpu.json
