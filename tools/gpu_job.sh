# one GPU-box visit: parity tests, bench (with and without the big unrolls), ncu launch list, ncu full captures
set -x; mkdir -p gpurun_out; T=${TAG:-v6}
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader; nproc
python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err || tail -20 gpurun_out/bench_$T.err
SCN_FAST_NO_BIG_UNROLL=1 python bench.py --steps 3 > gpurun_out/bench_${T}_nobig.json 2> gpurun_out/bench_${T}_nobig.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${T}_reference.json 2>/dev/null
python - <<PY
import json
for v in ("$T", "${T}_nobig"):
    d=json.loads([l for l in open('gpurun_out/bench_%s.json'%v) if l.startswith('{')][-1])
    print(v, round(d['value']/1e9,3), 'e2e', round(d['e2e']['value']/1e9,3), {k:round(x['evals_per_s']/1e9,3) for k,x in d['extra']['sweep'].items()}, d['extra']['parity_vs_cpu_max_rel_err'], d['extra'].get('medium_queries',{}).get('queries_per_s'), d['extra'].get('medium_queries_c5',{}).get('queries_per_s'))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv python bench.py --steps 2 --warmup 3 --e2e-points 4194304 > gpurun_out/ncu_launches_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fast_noise3d --launch-skip 3 -c 1 -f -o gpurun_out/fast_$T python bench.py --steps 1 --warmup 3 --no-extras --e2e-points 1048576 > gpurun_out/ncu_fast_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c1_$T python tools/run_queries.py c1_sphere_3d_none 20 1 > gpurun_out/ncu_march_c1_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_march --launch-skip 1 -c 1 -f -o gpurun_out/march_c5_$T python tools/run_queries.py c5_multires_nonstat_renewalplus 20 1 > gpurun_out/ncu_march_c5_$T.log 2>&1
ls -la gpurun_out
