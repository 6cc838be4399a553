#!/bin/bash
# tensor-core roofline sweep of the large-head forward (BASELINE configs[4]): prototypes x depth on a 1024x2048 map
for cfg in 256,64,100 512,128,100 1000,128,100 512,256,100 1000,256,100 1000,256,19 1024,256,112; do
  python tools/cfg5_bench.py --protos $cfg --reps 3 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('N,K,C=$cfg', 'kernel %.3f ms' % d['kernel_ms'], '%.0f TFLOP/s algorithmic' % d['algorithmic_tflops'], 'frac of TF32 peak %.3f' % d['frac_of_tf32_peak'], 'logits rel err %.1e' % d['logits_rel_err_vs_fp64_strip'])"
done
