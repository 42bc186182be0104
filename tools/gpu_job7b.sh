# full GPU job, part B: ncu captures of the three march kernels (kept under the 64 MiB copy-back limit)
set -x; mkdir -p gpurun_out; T=${TAG:-v30}
ncu --set full --clock-control none --import-source on -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c1_$T python tools/run_queries.py c1_sphere_3d_none 20 1 > gpurun_out/ncu_march_c1_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c5_$T python tools/run_queries.py c5_multires_nonstat_renewalplus 20 1 > gpurun_out/ncu_march_c5_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march1d --launch-skip 1 -c 1 -f -o gpurun_out/march1d_$T python tools/run_queries.py c3_1d_renewal 20 1 > gpurun_out/ncu_march1d_$T.log 2>&1
ls -la gpurun_out | tail -6
