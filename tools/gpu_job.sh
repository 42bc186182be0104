#!/bin/bash
# The one GPU job script (run under gpurun):   bash tools/gpu_job.sh <stage> [<stage> ...]      TAG=<suffix of the output files>
#   tests      pytest -m gpu + smoke()
#   bench      bench.py (both arms) at N=1
#   queries    queries/s of the BASELINE media (tools/run_queries.py)
#   launches   ncu launch list of the bench command (gpu__time_duration.sum)
#   ncu_fast | ncu_c1 | ncu_c5 | ncu_c3     one `ncu --set full` capture of that kernel
#   ncu_report condense the captures of this TAG into *_summary.txt on the box and delete them (they are too large to travel back)
#   sanitize   compute-sanitizer memcheck / racecheck / initcheck over tools/sanitize_workload.py
# Everything is written under gpurun_out/ (scratch); what is to be judged is copied into profiles/ afterwards.
set -x; mkdir -p gpurun_out; T=${TAG:-r02}
NCU="ncu --set full --clock-control none --import-source on"
for stage in "$@"; do case $stage in
tests)
  python -m pytest tests -m gpu -q -x > gpurun_out/pytest_$T.log 2>&1; (head -40; echo ...; tail -8) < gpurun_out/pytest_$T.log | cut -c1-240
  python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ;;
bench)
  python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err || tail -20 gpurun_out/bench_$T.err
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${T}_reference.json 2>/dev/null
  tail -c 3000 gpurun_out/bench_$T.json ;;
queries)
  for name in c1_sphere_3d_none c5_multires_nonstat_renewalplus c3_1d_renewal c4_1d_renewalplus_xy_mis; do
    python tools/run_queries.py $name 20 3 | tail -2; python tools/run_queries.py $name 20 3 secondary | tail -2
  done 2>&1 | tee gpurun_out/queries_$T.log ;;
launches)
  ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$T.csv python bench.py --steps 2 --warmup 3 --e2e-points 4194304 --render-scale 0.03 > gpurun_out/ncu_launches_$T.log 2>&1 ;;
ncu_fast)
  $NCU -k regex:k_fast_noise3d --launch-skip 2 -c 1 -f -o gpurun_out/fast_$T python bench.py --steps 2 --warmup 3 --no-extras --e2e-points 4194304 > gpurun_out/ncu_fast_$T.log 2>&1 ;;
ncu_c1)
  $NCU -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c1_$T python tools/run_queries.py c1_sphere_3d_none 20 1 > gpurun_out/ncu_march_c1_$T.log 2>&1 ;;
ncu_c5)
  $NCU -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c5_$T python tools/run_queries.py c5_multires_nonstat_renewalplus 20 1 > gpurun_out/ncu_march_c5_$T.log 2>&1 ;;
ncu_c3)
  $NCU -k regex:k_march1d --launch-skip 1 -c 1 -f -o gpurun_out/march1d_$T python tools/run_queries.py c3_1d_renewal 20 1 > gpurun_out/ncu_march1d_$T.log 2>&1 ;;
ncu_report)
  # condense every capture ON THE BOX (four .ncu-rep files exceed the 64 MiB that travel back) and drop the captures
  for rep in gpurun_out/*_$T.ncu-rep; do
    out=${rep%.ncu-rep}_summary.txt
    { python tools/ncu_summary.py $rep; python tools/ncu_brief.py $rep; python tools/ncu_footprint.py $rep; python tools/ncu_blocks.py $rep 1.0; python tools/ncu_lines.py $rep 0.8 3e6; } > $out 2>&1
    rm -f $rep
  done ;;
sanitize)
  for tool in memcheck racecheck initcheck; do
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_workload.py 13 > gpurun_out/sanitize_${tool}_$T.log 2>&1
    tail -4 gpurun_out/sanitize_${tool}_$T.log
  done ;;
*) echo "unknown stage $stage" ;;
esac; done
ls -la gpurun_out | tail -8
