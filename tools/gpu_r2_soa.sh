set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -12
run() {
  local label="$1"; shift
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 200 python bench.py --no-cpu --no-e2e --no-sweep --steps 50 --warmup 10 "$@" > gpurun_out/tp.json 2> gpurun_out/tp.err
  python - "$label" <<'PY'
import json,sys
try:
    d=json.loads(open('gpurun_out/tp.json').read().strip().splitlines()[-1])
    print('%-44s %8.2f us  frac %.3f' % (sys.argv[1], d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open('gpurun_out/tp.err').read()[-1500:])
PY
}
for n in 4096 65536 262144 1048576 4194304; do run "h soa $n" X=1 -- --envs $n --layout soa; done
run "h soa 1M nostats" X=1 -- --envs 1048576 --layout soa --no-stats
run "h soa 1M table global" DM_TABLE_SMEM=0 -- --envs 1048576 --layout soa
for n in 262144 1048576; do run "g1c soa $n" X=1 -- --robot unitree_g1 --combined --envs $n --layout soa; done
run "g1 walk soa 1M" X=1 -- --robot unitree_g1 --envs 1048576 --layout soa
