#!/bin/bash
# ring-depth sweep of the large-head forward at K = 128 (A tile 64 KB leaves room for 16 slots) and K = 256
for sl in 2 4 8 12 16; do
  echo "K=128 SLOTS=$sl: $(ADLM_BIG_SLOTS=$sl timeout 120 python tools/cfg5_bench.py --reps 3 --hw 512x1024 --protos 1000,128,100 2>&1 | grep -o '"kernel_ms": [0-9.]*')"
done
for sl in 2 4 8; do
  echo "K=256 SLOTS=$sl: $(ADLM_BIG_SLOTS=$sl timeout 120 python tools/cfg5_bench.py --reps 3 --hw 512x1024 2>&1 | grep -o '"kernel_ms": [0-9.]*')"
done
