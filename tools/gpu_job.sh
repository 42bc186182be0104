set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for m in c1_sphere_3d_none c5_multires_nonstat_renewalplus; do timeout 300 python tools/run_queries.py $m 20 2; done > gpurun_out/queries_all.log 2>&1
cat gpurun_out/queries_all.log
Q="python tools/run_queries.py c5_multires_nonstat_renewalplus 18 2"
$Q > gpurun_out/queries_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_march -s 1 -c 1 -o gpurun_out/prof_march_c5 $Q > gpurun_out/ncu3.log 2>&1
