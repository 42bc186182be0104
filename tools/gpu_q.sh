# developer job: GPU parity tests, then query throughput of the BASELINE media (tools/run_queries.py)
set -x; mkdir -p gpurun_out; T=${TAG:-q}
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
for name in c1_sphere_3d_none c5_multires_nonstat_renewalplus; do
  python tools/run_queries.py $name 20 3 | tail -2
  python tools/run_queries.py $name 20 3 secondary | tail -2
done 2>&1 | tee gpurun_out/queries_$T.log
