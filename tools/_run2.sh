set -x
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_minpool.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -15 > $O/r1_pytest_minpool.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 > $O/r1_pytest_gpu.log
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
(echo '# python tools/minpool_bench.py 16 (B200)'; timeout 300 python tools/minpool_bench.py 16) > $O/r1_min_pool.log 2>&1
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
cat $O/r1_pytest_minpool.log $O/r1_pytest_gpu.log $O/r1_kernel_sweep_tf32.log $O/r1_min_pool.log
ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS timeout 300 python -m adlm_inherently_interpretable_segmentation_b200.build --force
(timeout 100 python tools/phase_clocks_ws.py 16) > $O/phase_ws.log 2>&1
(timeout 100 python tools/phase_clocks_dp.py 16) > $O/phase_dp.log 2>&1
cat $O/phase_ws.log $O/phase_dp.log
