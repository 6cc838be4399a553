timeout 300 python -m pytest tests/test_gpu_head.py -m gpu -q -k 57_tiles 2>&1 | grep -E "^E|assert|passed|failed" | head -20 > gpurun_out/t57.log; cat gpurun_out/t57.log
