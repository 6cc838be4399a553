set -x
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_step.py -m gpu -q 2>&1 | tail -3 > $O/pytest_step.log
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
cat $O/pytest_step.log $O/r1_kernel_sweep_tf32.log
