set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
for m in c1_sphere_3d_none c5_multires_nonstat_renewalplus multires_world; do timeout 300 python tools/run_queries.py $m 20 2; timeout 300 python tools/run_queries.py $m 20 2 secondary; done > gpurun_out/queries_all.log 2>&1
cat gpurun_out/queries_all.log
