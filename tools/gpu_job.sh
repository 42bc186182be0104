set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
for c in c1 c5; do timeout 600 python tools/render.py --config $c --res 256 --spp 16 --out gpurun_out/$c.pfm; timeout 600 python tools/render.py --config $c --res 256 --spp 16 --count-queries; done > gpurun_out/render.log 2>&1
cat gpurun_out/render.log
for m in c5_multires_nonstat_renewalplus; do timeout 300 python tools/run_queries.py $m 20 2; done
