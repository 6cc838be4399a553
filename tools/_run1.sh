set -x
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_minpool.py -m gpu -x -q 2>&1 | tail -15 > $O/r1_pytest_minpool.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 > $O/r1_pytest_gpu.log
(echo '# python tools/nearest_bench.py 64 (B200)'; timeout 300 python tools/nearest_bench.py 64) > $O/r1_k_nearest.log 2>&1
(echo '# python tools/minpool_bench.py 16 (B200)'; timeout 300 python tools/minpool_bench.py 16) > $O/r1_min_pool.log 2>&1
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
cat $O/r1_pytest_minpool.log $O/r1_pytest_gpu.log $O/r1_k_nearest.log $O/r1_min_pool.log $O/r1_bench_tf32.json
