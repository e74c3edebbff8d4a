P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print("kernel_ms", round(d["roofline"]["kernel_ms"],5), "frac", round(d["roofline"]["frac"],4))'
for n in 4096 32768 65536 262144 1048576; do for v in "DM_ROW_STAGE=0" "DM_ROW_STAGE=1" "DM_ROW_STAGE=1 DM_WIDE=2"; do
  echo -n "G1 combined envs=$n $v: "; env $v python bench.py --robot unitree_g1 --combined --envs $n --steps 100 --warmup 10 --no-cpu --no-e2e --no-sweep 2>/dev/null | python -c "$P"
done; done
