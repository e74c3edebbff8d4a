# final single-GPU pass: GPU tests, default bench line (with the CPU arm), reference arm, ncu captures of every config shape
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02_bench_1gpu_final.json 2> gpurun_out/r02_bench_1gpu_final.err; tail -c 600 gpurun_out/r02_bench_1gpu_final.json; echo
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/ref.err; tail -c 400 gpurun_out/r02_bench_reference_arm.json; echo
bash tools/gpu_r2_ncu_all.sh 2>&1 | tail -14
