set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -q -k "microbench or golden" 2>&1 | tail -2
SCN_LIB_PATH=$PWD/sparse_conv_gpis_tungsten_b200/libscn_var0.so python bench.py --steps 3 --warmup 3 > gpurun_out/bench_var0.json 2>gpurun_out/bench_var0.err
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_var2.json 2>gpurun_out/bench_var2.err
python - <<'PY'
import json
for v in (0,2):
    d=json.loads([l for l in open('gpurun_out/bench_var%d.json'%v) if l.startswith('{')][-1])
    print(v, round(d['value']/1e9,3), {k:round(x['evals_per_s']/1e9,3) for k,x in d['extra']['sweep'].items() if k in ('iso_N1','iso_N2','iso_N4','iso_N8','world_N1','world_N4')}, d['extra']['parity_vs_cpu_max_rel_err'])
PY
