set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
for m in c1_sphere_3d_none world_3d_none iso_3d_renewalplus_aniso; do timeout 300 python tools/run_queries.py $m 20 2; timeout 300 python tools/run_queries.py $m 20 2 secondary; done > gpurun_out/queries_all.log 2>&1
cat gpurun_out/queries_all.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -3 gpurun_out/bench_full.err
