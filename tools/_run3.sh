set -x
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_minpool.py tests/test_gpu_push.py tests/test_gpu_step.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -15 > $O/r1_pytest_minpool.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 > $O/r1_pytest_gpu.log
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
cat $O/r1_pytest_minpool.log $O/r1_pytest_gpu.log $O/r1_kernel_sweep_tf32.log
