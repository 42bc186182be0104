set -x; mkdir -p gpurun_out
N=${NG:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -2 gpurun_out/bench_n1.err
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -3 gpurun_out/bench_n$N.err
timeout 600 python tools/render.py --config c1 --res 1024 --spp 16 > gpurun_out/render_scale.log 2>&1
timeout 600 $TR tools/render.py --config c1 --res 1024 --spp 16 >> gpurun_out/render_scale.log 2>&1
timeout 600 python tools/render.py --config c5 --res 1024 --spp 16 >> gpurun_out/render_scale.log 2>&1
timeout 600 $TR tools/render.py --config c5 --res 1024 --spp 16 >> gpurun_out/render_scale.log 2>&1
grep workload gpurun_out/render_scale.log
