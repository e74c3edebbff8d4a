set -u
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-sweep"
name=h2M; rx=step_tile
$B --envs 2097152 > gpurun_out/plain_$name.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$name.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1 -f -o gpurun_out/r02_$name $B --envs 2097152 > gpurun_out/ncu_$name.log 2>&1
{ echo "# $B --envs 2097152"; echo "# ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 1"
  python tools/ncu_summary.py gpurun_out/r02_$name.ncu-rep
  echo; echo "# per-instruction stall samples (SASS view)"; python tools/ncu_hot.py gpurun_out/r02_$name.ncu-rep stall_long_sb 12; } > gpurun_out/r02_${name}_ncu_full_summary.txt 2>&1
rm -f gpurun_out/r02_$name.ncu-rep
grep -E "gpu__time_duration|dram__bytes|issue_active|registers_per_thread" gpurun_out/r02_${name}_ncu_full_summary.txt
