set -x; mkdir -p gpurun_out; T=${TAG:-v9}
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
for name in c3_1d_renewal c4_1d_renewalplus_xy_mis 1d_none_single; do
  for mode in primary secondary; do
    python tools/run_queries.py $name 20 3 $mode | tail -2
    SCN_FORCE_GENERIC=1 python tools/run_queries.py $name 20 3 $mode | tail -2
  done
done > gpurun_out/ab1d_$T.log 2>&1; grep rays gpurun_out/ab1d_$T.log
