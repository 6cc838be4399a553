O=gpurun_out
mkdir -p $O
L=adlm_inherently_interpretable_segmentation_b200/libadlm_head.so
cp $L /tmp/keep.so
for v in new old new old; do
  cp tools/_bin/$v.so $L
  echo "== variant $v"
  timeout 120 python tools/kernel_sweep.py --batches 256 2>&1 | grep -E "B= 256|Error" | head -8
done > $O/ab_reduce.log 2>&1
cp /tmp/keep.so $L
cat $O/ab_reduce.log
