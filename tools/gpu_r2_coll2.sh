set -u
mkdir -p gpurun_out
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    c=d['config'].get('collective')
    print('%-40s n=%d  %8.2f us/step  kernel %.2f us  coll=%s count=%s/%s launches=%s' % (sys.argv[1], d['n_gpus'], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3,
          (c or {}).get('mode'), d.get('stats',{}).get('count'), d.get('stats',{}).get('count_expected'), d['gpu_launches']))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[2].replace('.json','.err')).read()[-1200:])
PY
}
Q="--no-cpu --no-e2e --no-sweep --steps 200 --warmup 20"
for dbg in 2 3 1; do
  DM_PEER_DEBUG=$dbg timeout 200 python bench.py $Q --collective auto --collective-at-1 > gpurun_out/d_$dbg.json 2> gpurun_out/d_$dbg.err; show "N=1 auto DM_PEER_DEBUG=$dbg" gpurun_out/d_$dbg.json
done
DM_PDL=0 timeout 200 python bench.py $Q --collective auto --collective-at-1 > gpurun_out/d_nopdl.json 2> gpurun_out/d_nopdl.err; show "N=1 auto DM_PDL=0" gpurun_out/d_nopdl.json
DM_PDL=0 timeout 200 python bench.py $Q --collective off > gpurun_out/d_nopdl0.json 2> gpurun_out/d_nopdl0.err; show "N=1 off DM_PDL=0" gpurun_out/d_nopdl0.json
timeout 200 python bench.py $Q --collective auto --collective-at-1 --envs 65536 > gpurun_out/d_64k.json 2> gpurun_out/d_64k.err; show "N=1 auto 65536" gpurun_out/d_64k.json
timeout 200 python bench.py $Q --collective off --envs 65536 > gpurun_out/d_64k0.json 2> gpurun_out/d_64k0.err; show "N=1 off 65536" gpurun_out/d_64k0.json
