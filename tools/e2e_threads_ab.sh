# e2e of bench.py (DPVecEnv.step, pageable inputs, 4096 envs) against the number of copy helper threads, three runs each
P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); e=d["e2e"]; print("%.1f us" % (e["ms_per_step"]*1e3), end="  ")'
for t in 0 1 2 3 5; do echo -n "DM_COPY_THREADS=$t: "; for r in 1 2 3; do DM_COPY_THREADS=$t python bench.py --no-cpu --no-sweep --steps 200 --warmup 20 2>/dev/null | python -c "$P"; done; echo; done
