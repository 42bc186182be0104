set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -q -k "microbench or primitives or golden" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
tail -4 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_full.err | tail -5
