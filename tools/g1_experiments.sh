B="python bench.py --robot unitree_g1 --combined --envs 1048576 --steps 20 --warmup 5 --no-cpu --no-e2e --no-sweep"
P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d["roofline"]["kernel_ms"], d["roofline"]["frac"])'
echo base; $B 2>/dev/null | python -c "$P"
echo nostats; $B --no-stats 2>/dev/null | python -c "$P"
echo skeleton; DM_SKIP_COMPUTE=1 $B 2>/dev/null | python -c "$P"
echo groups2; DM_MAX_GROUPS=2 $B 2>/dev/null | python -c "$P"
echo g1-dpenv-walk; python bench.py --robot unitree_g1 --motion walk --envs 1048576 --steps 20 --warmup 5 --no-cpu --no-e2e --no-sweep 2>/dev/null | python -c "$P"
echo g1-262144; python bench.py --robot unitree_g1 --combined --envs 262144 --steps 50 --warmup 5 --no-cpu --no-e2e --no-sweep 2>/dev/null | python -c "$P"
