set -x; mkdir -p gpurun_out; T=${TAG:-v8}
timeout 600 python -m pytest tests/test_wavefront.py -m gpu -q -s 2>&1 | grep -E "relMSE|passed|failed|Error|assert|error" | head -60 > gpurun_out/pytest_$T.log; cat gpurun_out/pytest_$T.log
