set -x; mkdir -p gpurun_out; T=${TAG:-v7}
python -m pytest tests -m gpu -q -x -s -k "images_vs_reference or golden or wavefront" 2>&1 | grep -E "relMSE|passed|failed|Error|assert" | head -40 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
for name in c1_sphere_3d_none c5_multires_nonstat_renewalplus; do
  python tools/run_queries.py $name 20 3 | tail -2
  SCN_LIB_PATH=$PWD/sparse_conv_gpis_tungsten_b200/libscn_var0.so python tools/run_queries.py $name 20 3 | tail -2
  python tools/run_queries.py $name 20 3 secondary | tail -2
  SCN_LIB_PATH=$PWD/sparse_conv_gpis_tungsten_b200/libscn_var0.so python tools/run_queries.py $name 20 3 secondary | tail -2
done > gpurun_out/ab_$T.log 2>&1; cat gpurun_out/ab_$T.log
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
