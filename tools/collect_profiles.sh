set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/r1_pytest_gpu.log
timeout 300 python bench.py > $O/r1_bench_tf32.json 2>$O/r1_bench_tf32.err
timeout 300 python bench.py --precision tf32x3 --no-cpu > $O/r1_bench_tf32x3.json 2>$O/err2
timeout 300 python bench.py --precision fp32 --no-cpu > $O/r1_bench_fp32.json 2>$O/err3
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > $O/r1_bench_reference.json 2>$O/err4
timeout 200 python tools/kernel_sweep.py --batches 16,64,256 > $O/r1_kernel_sweep_tf32.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r1_launches_tf32_step.csv python bench.py --steps 10 --warmup 3 --no-cpu --no-push > $O/ncu_l.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'k_tc_|k_kld|k_step_prepare|k_reduce' -c 12 -o $O/r1_step_full -f python tools/kernel_sweep.py --once --batches 16 > $O/ncu_f.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'k_tc_push|k_patch_class_mask' -c 4 -o $O/r1_push_full -f python tools/push_once.py 64 tf32x3 > $O/ncu_p.log 2>&1
timeout 200 python tools/sigmoid_fusion_bench.py > $O/r1_sigmoid_fusion.log 2>&1
(echo '# python tools/eval_bench.py (B200, 4 images, CUDA events)'; timeout 300 python tools/eval_bench.py) > $O/r1_eval_upsample.log 2>&1
(echo '# python tools/nearest_bench.py 64 (B200)'; timeout 300 python tools/nearest_bench.py 64) > $O/r1_k_nearest.log 2>&1
(echo '# python tools/minpool_bench.py 16 (B200)'; timeout 300 python tools/minpool_bench.py 16) > $O/r1_min_pool.log 2>&1
ls -la $O | tail -24
