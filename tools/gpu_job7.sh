# full GPU job, part A: parity tests, smoke, bench (both arms), launch list, ncu capture of the roofline kernel
set -x; mkdir -p gpurun_out; T=${TAG:-v30}
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err || tail -20 gpurun_out/bench_$T.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${T}_reference.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_$T.json') if l.startswith('{')][-1])
print(round(d['value']/1e9,3), 'e2e', round(d['e2e']['value']/1e9,3), 'frac', round(d['roofline']['frac'],3), 'traffic', d['roofline']['traffic'], 'cpu', d['cpu_baseline'])
for k,v in d['extra'].items():
    if k.startswith('medium_queries') or k.startswith('render'): print(k, json.dumps(v)[:400])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$T.csv python bench.py --steps 2 --warmup 3 --e2e-points 4194304 > gpurun_out/ncu_launches_$T.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fast_noise3d --launch-skip 2 -c 1 -f -o gpurun_out/fast_$T python bench.py --steps 2 --warmup 3 --no-extras --e2e-points 4194304 > gpurun_out/ncu_fast_$T.log 2>&1
ls -la gpurun_out | tail -6
