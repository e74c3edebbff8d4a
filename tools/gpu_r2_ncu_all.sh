# ncu captures of this round (roofline.traffic from a capture of each config's very shape).
# Every command runs WITHOUT ncu first (numbers printed under a profiler are never bench values).
set -u
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep"
cap() {  # name, kernel regex, bench args...
  local name="$1" rx="$2"; shift 2
  $B "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "$name: plain run failed"; tail -5 gpurun_out/plain_$name.log; return; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1 -f -o gpurun_out/r02_$name $B "$@" > gpurun_out/ncu_$name.log 2>&1
  # the reports together exceed gpurun's 64 MiB return limit: summarise here, keep the text
  { echo "# $B $*"; echo "# ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1"
    python tools/ncu_summary.py gpurun_out/r02_$name.ncu-rep
    echo; echo "# per-instruction stall samples (SASS view)"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_long_sb 12
    echo "--- top stall_no_inst"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_no_inst 6 | tail -7; } > gpurun_out/r02_${name}_ncu_full_summary.txt 2>&1
  rm -f gpurun_out/r02_$name.ncu-rep
  grep -E "gpu__time_duration|dram__bytes" gpurun_out/r02_${name}_ncu_full_summary.txt | awk -v n=$name '{printf "%s %s %s %s | ", n, $1, $2, $3} END {print ""}'
}
cap h4096 step_tile --envs 4096
cap c3_8192 step_tile --envs 8192 --motion spinkick,dance_b
cap c3_65536 step_tile --envs 65536 --motion spinkick,dance_b
cap h1M step_tile --envs 1048576
cap g1c_32768 step_tile --robot unitree_g1 --combined --envs 32768
cap g1c_262144 step_tile --robot unitree_g1 --combined --envs 262144
cap g1c_1M step_tile --robot unitree_g1 --combined --envs 1048576
cap soa_h1M step_soa --envs 1048576 --layout soa
cap soa_g1c_1M step_soa --robot unitree_g1 --combined --envs 1048576 --layout soa
# launch list of the default bench command (the driver's), without the CPU leg
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain_default.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r02_launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_default.log 2>&1
python tools/launch_summary.py gpurun_out/r02_launches_default.csv > gpurun_out/r02_launches_default_summary.txt; head -12 gpurun_out/r02_launches_default_summary.txt
