set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for v in -1 1; do
  echo "== variant $v humanoid"; timeout 300 python bench.py --no-cpu --no-e2e --variant $v --sweep-envs 32768,65536,262144,1048576 > gpurun_out/bench_h_v$v.json 2> gpurun_out/bench_h_v$v.err; tail -c 1500 gpurun_out/bench_h_v$v.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]) if False else None
" 2>/dev/null; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_h_v$v.json').read().strip().splitlines()[-1])
    print('4096:', d['roofline']['kernel_ms']*1e3, 'us frac', d['roofline']['frac'])
    for r in d.get('sweep',[]): print(r['envs'], round(r['kernel_ms']*1e3,2), 'us frac', round(r['frac'],3))
except Exception as e: print('ERR', e); print(open('gpurun_out/bench_h_v$v.err').read()[-2000:])
PY
  echo "== variant $v g1 combined"; timeout 300 python bench.py --no-cpu --no-e2e --variant $v --robot unitree_g1 --combined --envs 32768 --sweep-envs 262144,1048576 > gpurun_out/bench_c_v$v.json 2> gpurun_out/bench_c_v$v.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_c_v$v.json').read().strip().splitlines()[-1])
    print('32768:', d['roofline']['kernel_ms']*1e3, 'us frac', d['roofline']['frac'])
    for r in d.get('sweep',[]): print(r['envs'], round(r['kernel_ms']*1e3,2), 'us frac', round(r['frac'],3))
except Exception as e: print('ERR', e); print(open('gpurun_out/bench_c_v$v.err').read()[-2000:])
PY
done
