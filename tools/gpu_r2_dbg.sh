set -u
mkdir -p gpurun_out
R="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5 --envs 1048576"
for v in A D E C; do
  cp _variants/$v.so deepmimic_mujoco_b200/libdmhot.so
  timeout 300 $R --steps 60 > gpurun_out/v$v.json 2> gpurun_out/v$v.err; echo "variant $v steps=60 rc=$? $(grep -o 'line [0-9]*, in [a-z_]*' gpurun_out/v$v.err | tail -2 | tr '\n' ' ')"
done
cp _variants/A.so deepmimic_mujoco_b200/libdmhot.so
timeout 300 $R --steps 12 > gpurun_out/vA6.json 2> gpurun_out/vA6.err; echo "A steps=12 rc=$? $(grep -o 'line [0-9]*, in [a-z_]*' gpurun_out/vA6.err | tail -2 | tr '\n' ' ')"
DM_PDL=0 timeout 300 $R --steps 60 > gpurun_out/vA7.json 2> gpurun_out/vA7.err; echo "A PDL=0 steps=60 rc=$?"
DM_NO_PREFETCH=1 timeout 300 $R --steps 60 > gpurun_out/vA8.json 2> gpurun_out/vA8.err; echo "A no first-tile prefetch steps=60 rc=$?"
