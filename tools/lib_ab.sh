# kernel-variant experiment: bench the step kernel with an alternative libdmhot.so (argument) swapped in on the GPU box
cp "$1" deepmimic_mujoco_b200/libdmhot.so
P='import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d["config"]["workload"][:44], "kernel_ms", round(d["roofline"]["kernel_ms"],5), "frac", round(d["roofline"]["frac"],4))'
for n in ${SIZES:-4096 65536 262144 1048576}; do
  python bench.py --envs $n --steps 100 --warmup 10 --no-cpu --no-e2e --no-sweep $EXTRA 2>/dev/null | python -c "$P"
done
