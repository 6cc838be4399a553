O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/r1_pytest_gpu.log
for i in 1 2 3 4; do echo "== plain $i"; timeout 120 python tools/kernel_sweep.py --batches 256 2>&1 | grep -E "B= 256 M|Error|adlm" | head -8; done > $O/repro3.log 2>&1
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
cat $O/r1_pytest_gpu.log $O/repro3.log $O/r1_kernel_sweep_tf32.log; head -c 300 $O/r1_bench_tf32.json
