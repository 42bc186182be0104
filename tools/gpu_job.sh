set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1
B="python bench.py --steps 2 --warmup 3 --points 16777216 --e2e-points 1048576 --no-extras"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_fast_noise3d -s 3 -c 2 -o gpurun_out/prof_fast $B > gpurun_out/ncu2.log 2>&1
Q="python tools/run_queries.py c1_sphere_3d_none 18 2"
$Q > gpurun_out/queries_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_sample_distance -s 1 -c 1 -o gpurun_out/prof_march $Q > gpurun_out/ncu3.log 2>&1
for m in c1_sphere_3d_none c5_multires_nonstat_renewalplus c3_1d_renewal world_3d_none; do python tools/run_queries.py $m 20 2; python tools/run_queries.py $m 20 2 secondary; done > gpurun_out/queries_all.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/queries_all.log
