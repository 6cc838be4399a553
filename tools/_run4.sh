set -x
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/r1_pytest_gpu.log
timeout 200 python tools/kernel_sweep.py --batches 16,64 > $O/r1_kernel_sweep_tf32.log 2>&1
cat $O/r1_pytest_gpu.log $O/r1_kernel_sweep_tf32.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'k_kld|k_step_prepare' -c 2 -o $O/r1_kld_full -f python tools/kernel_sweep.py --once --batches 16 > $O/ncu_kld.log 2>&1
tail -3 $O/ncu_kld.log
