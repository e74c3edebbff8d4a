# last pass of round 2: the soak tests, then the driver's default bench line (refreshes profiles/r02_bench_1gpu_final.json)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_soak.py tests/test_gpu_batch.py -x -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02_bench_1gpu_final.json 2> gpurun_out/r02_bench_1gpu_final.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_1gpu_final.json; echo
