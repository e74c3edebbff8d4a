set -u
mkdir -p gpurun_out
one() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-40s %8.2f us/step kernel %.2f us frac %.3f' % (sys.argv[1], d['ms_per_step']*1e3, d['roofline']['kernel_ms']*1e3, d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
}
R1="python bench.py --gpus 1 --no-cpu --no-e2e --no-sweep --no-configs --warmup 5 --steps 30"
for rep in 1 2; do
for pn in 0 1; do
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs 1048576 --layout soa > gpurun_out/g.json 2>gpurun_out/g.err; one "h soa envs=1M next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs 262144 --layout soa > gpurun_out/g.json 2>gpurun_out/g.err; one "h soa envs=262144 next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs 1048576 --layout soa --robot unitree_g1 > gpurun_out/g.json 2>gpurun_out/g.err; one "g1 DPEnv soa envs=1M next=$pn" gpurun_out/g.json
DM_PREFETCH_NEXT=$pn timeout 300 $R1 --envs 1048576 --layout soa --robot unitree_g1 --combined > gpurun_out/g.json 2>gpurun_out/g.err; one "g1c soa envs=1M next=$pn" gpurun_out/g.json
done
done
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep"
cap() {  # name, kernel regex, bench args...
  local name="$1" rx="$2"; shift 2
  $B "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "$name: plain run failed"; tail -5 gpurun_out/plain_$name.log; return; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1 -f -o gpurun_out/r02_$name $B "$@" > gpurun_out/ncu_$name.log 2>&1
  { echo "# $B $*"; echo "# ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1"
    python tools/ncu_summary.py gpurun_out/r02_$name.ncu-rep
    echo; echo "# per-instruction stall samples (SASS view)"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_long_sb 12
    echo "--- top stall_short_sb"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_short_sb 8 | tail -9
    echo "--- top stall_barrier"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_barrier 6 | tail -7
    echo "--- top stall_no_inst"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_no_inst 6 | tail -7; } > gpurun_out/r02b_${name}_ncu_full_summary.txt 2>&1
  rm -f gpurun_out/r02_$name.ncu-rep
}
cap soa_g1c_1M step_soa --robot unitree_g1 --combined --envs 1048576 --layout soa
