set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-44s %8.2f us/step kernel %.2f us frac %.3f' % (sys.argv[1], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
    for k,v in (d.get('configs') or {}).items():
        if isinstance(v,dict): print('     ',k,{a:b for a,b in v.items() if a in('envs_per_gpu','ms_per_step','kernel_ms','frac')})
        else:
            for r in v: print('      ', r['envs_per_gpu'], {a:(round(b['kernel_ms']*1e3,1), round(b['frac'],3)) for a,b in r.items() if isinstance(b,dict)})
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --warmup 5"
timeout 600 $R1 --steps 20 > gpurun_out/full1.json 2>gpurun_out/g.err; one "default N=1 with configs" gpurun_out/full1.json
timeout 600 $R1 --steps 200 --no-configs --no-sweep > gpurun_out/g.json 2>gpurun_out/g.err; one "N=1 200 steps" gpurun_out/g.json
