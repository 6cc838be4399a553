set -x
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/r1_pytest_gpu.log
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
(echo "# python tools/kernel_sweep.py --protos 30,3 --infer --batches 64,256,1024 (cfg4: mds_new pancreas head, inference + per-pixel argmax)"; timeout 200 python tools/kernel_sweep.py --protos 30,3 --infer --batches 64,256,1024) > $O/r1_kernel_sweep_cfg4.log 2>&1
cat $O/r1_pytest_gpu.log $O/r1_kernel_sweep_tf32.log $O/r1_kernel_sweep_cfg4.log
