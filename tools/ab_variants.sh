#!/bin/bash
# Developer A/B of library builds (SCN_LIB_PATH): eval rate (bench.py --no-extras) and C1 / C5 / C3 query rates per build.
#   bash tools/ab_variants.sh libscn_b200.so libscn_b200_v1.so ...
mkdir -p gpurun_out
for lib in "$@"; do
  export SCN_LIB_PATH=$PWD/sparse_conv_gpis_tungsten_b200/$lib
  echo "== $lib"
  python bench.py --no-extras --steps 20 --warmup 3 --e2e-points 1048576 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('eval %.4g %s  ms/step %.3f' % (d['value'], d['unit'], d['ms_per_step']))"
  for m in c1_sphere_3d_none c5_multires_nonstat_renewalplus c3_1d_renewal; do python tools/run_queries.py $m 20 3 2>&1 | grep "queries/s" | tail -1 | cut -c1-120; done
done 2>&1 | tee gpurun_out/ab_variants.log
