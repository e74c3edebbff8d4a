# programmatic dependent launch on / off: kernel time per step from bench.py's CUDA-event region
P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d["config"]["workload"][:40], "ms/step", round(d["ms_per_step"],5), "kernel_ms", round(d["roofline"]["kernel_ms"],5), "frac", round(d["roofline"]["frac"],4))'
for pdl in 0 1; do for n in 4096 16384 65536 1048576; do
  echo -n "DM_PDL=$pdl envs=$n: "; DM_PDL=$pdl python bench.py --envs $n --steps 200 --warmup 20 --no-cpu --no-e2e --no-sweep 2>/dev/null | python -c "$P"
done; done
