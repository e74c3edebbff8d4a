set -u
mkdir -p gpurun_out
echo "== alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15
echo "== whole gpu suite"; timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
echo "== sanitizer on the new entry point"; timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/sanitize_noalloc.log 2>&1; tail -6 gpurun_out/sanitize_noalloc.log
python bench.py --no-cpu --no-e2e --steps 50 --warmup 10 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -c 600 gpurun_out/bench_quick.json
