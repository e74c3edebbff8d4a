set -u
mkdir -p gpurun_out
run() {  # label, env assignments..., -- bench args
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
  grep "dmhot" gpurun_out/tp.err | head -24
}
timeout 300 python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py tests/test_gpu_batch.py -x -q -m gpu 2>&1 | tail -5
echo "=== humanoid 4096"
run "h4096 team trace" DM_TEAM_TRACE=1 -- --envs 4096
run "h4096 team nostats trace" DM_TEAM_TRACE=1 -- --envs 4096 --no-stats
run "h4096 team skip_compute" DM_SKIP_COMPUTE=1 -- --envs 4096
run "h4096 team table=0" DM_TEAM_TABLE=0 -- --envs 4096
run "h4096 team nopdl" DM_PDL=0 -- --envs 4096
run "h4096 tile" X=1 -- --envs 4096 --variant 1
echo "=== humanoid 1M"
run "h1M team auto" DM_TEAM_TRACE=1 -- --envs 1048576
run "h1M team nin2" DM_TEAM_NIN=2 -- --envs 1048576
run "h1M team maxt6" DM_TEAM_MAXT=6 -- --envs 1048576
run "h1M team maxt6 nin2" DM_TEAM_MAXT=6 DM_TEAM_NIN=2 -- --envs 1048576
run "h1M team maxt6 table1" DM_TEAM_MAXT=6 DM_TEAM_TABLE=1 -- --envs 1048576
run "h1M team maxt4 nin2" DM_TEAM_MAXT=4 DM_TEAM_NIN=2 -- --envs 1048576
run "h1M team nostats" X=1 -- --envs 1048576 --no-stats
run "h1M team skip" DM_SKIP_COMPUTE=1 -- --envs 1048576
echo "=== g1 combined 262144"
run "c262k team auto" DM_TEAM_TRACE=1 -- --robot unitree_g1 --combined --envs 262144
run "c262k team nin2" DM_TEAM_NIN=2 -- --robot unitree_g1 --combined --envs 262144
run "c262k team maxt6" DM_TEAM_MAXT=6 -- --robot unitree_g1 --combined --envs 262144
run "c262k team maxt4 nin2" DM_TEAM_MAXT=4 DM_TEAM_NIN=2 -- --robot unitree_g1 --combined --envs 262144
run "c262k team skip" DM_SKIP_COMPUTE=1 -- --robot unitree_g1 --combined --envs 262144
run "c262k tile" X=1 -- --robot unitree_g1 --combined --envs 262144 --variant 1
