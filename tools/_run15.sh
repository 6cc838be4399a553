O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/r1_pytest_gpu.log
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1
cat $O/r1_pytest_gpu.log $O/r1_kernel_sweep_tf32.log $O/smoke.log; head -c 250 $O/r1_bench_tf32.json
