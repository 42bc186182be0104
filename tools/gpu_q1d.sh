# developer job: GPU parity tests, then query throughput of the 1D BASELINE media (tools/run_queries.py)
set -x; mkdir -p gpurun_out; T=${TAG:-q1d}
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
for name in c3_1d_renewal c4_1d_renewalplus_xy_mis; do
  python tools/run_queries.py $name 20 3 | tail -2
  python tools/run_queries.py $name 20 3 secondary | tail -2
done 2>&1 | tee gpurun_out/queries_$T.log
